"""Opcode mix and stall samples of one kernel from an ncu source-page CSV:
    ncu -i x.ncu-rep --page source --csv > x_src.csv ; python tools/ncu_opmix.py x_src.csv [elements]
"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
ops = collections.defaultdict(lambda: [0, 0, 0])
tot_i = tot_s = 0
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    src = r[col["Source"]].strip()
    toks = src.split()
    if not toks:
        continue
    op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
    op = op.rstrip(";")
    base = op.split(".")[0]
    n = int(r[col["Instructions Executed"]] or 0)
    s = int(r[col["# Samples"]] or 0)
    ops[base][0] += n
    ops[base][1] += s
    ops[base][2] += 1
    tot_i += n
    tot_s += s
elems = float(sys.argv[2]) if len(sys.argv) > 2 else None
print(f"total warp instructions {tot_i:,}  stall samples {tot_s:,}" + (f"  thread-instr/element {tot_i * 32 / elems:.2f}" if elems else ""))
for k, (n, s, c) in sorted(ops.items(), key=lambda kv: -kv[1][0])[:40]:
    per = f"{n * 32 / elems:6.2f}/elt" if elems else ""
    print(f"{k:12s} {n:12,d} {100 * n / tot_i:5.1f}%  {per}  samples {100 * s / max(tot_s, 1):5.1f}%  sites {c}")
