"""Top source lines by warp-stall samples from an ncu report captured with --import-source on (-lineinfo build).
    python tools/ncu_hot.py gpurun_out/x.ncu-rep [N]"""
import csv
import subprocess
import sys

rep, n = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None
for i, r in enumerate(rows):
    if "Source" in r and any("Sampl" in c for c in r):
        hdr, rows = r, rows[i + 1:]
        break
if hdr is None:
    print(out[:2000])
    sys.exit(0)
ci = {h: i for i, h in enumerate(hdr)}
scol = [h for h in hdr if h.startswith("# Samples") or h == "Warp Stall Sampling (All Samples)"]
scol = scol[0] if scol else [h for h in hdr if "Sampl" in h][0]
tot = 0
items = []
for r in rows:
    try:
        v = float(r[ci[scol]])
    except Exception:
        continue
    tot += v
    items.append((v, r[ci["Source"]][:110]))
print("column:", scol, "total samples", tot)
for v, s in sorted(items, reverse=True)[:n]:
    print(f"{v:8.0f} {100 * v / max(tot, 1):5.1f}%  {s}")
