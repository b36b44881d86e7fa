"""Top stall sites (with the two leading stall reasons) per kernel of an ncu source-page CSV:
    ncu -i x.ncu-rep --page source --csv > x_src.csv ; python tools/ncu_stalls.py x_src.csv [kernel substring] [n]
"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2] if len(sys.argv) > 2 else ""
n = int(sys.argv[3]) if len(sys.argv) > 3 else 20
sections, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}
        sections.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None and len(r) >= len(cur["hdr"]):
        cur["rows"].append(r)
seen = set()
for s in sections:
    if want not in s["name"] or s["name"] in seen:
        continue
    seen.add(s["name"])
    col = {h: i for i, h in enumerate(s["hdr"])}
    sc = [h for h in s["hdr"] if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[col["# Samples"]] or 0) for r in s["rows"])
    print("====", s["name"][:120], "samples", tot)
    for r in sorted(s["rows"], key=lambda r: -int(r[col["# Samples"]] or 0))[:n]:
        st = sorted(((int(r[col[h]] or 0), h[6:]) for h in sc), reverse=True)[:2]
        print(f'{int(r[col["# Samples"]]):5d} {100 * int(r[col["# Samples"]]) / max(tot, 1):4.1f}% exec={int(r[col["Instructions Executed"]]):8d} '
              f'{r[col["Source"]].strip()[:84]:84s} {st}')
