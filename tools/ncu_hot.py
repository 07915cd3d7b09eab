#!/usr/bin/env python
"""Hottest SASS instructions of an `ncu --page source --csv` export (warp-stall samples per instruction).

    ncu -i report.ncu-rep --page source --csv > src.csv ; python tools/ncu_hot.py src.csv [N]
"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
col = hdr.index("Warp Stall Sampling (All Samples)")
body = [r for r in rows[h + 1:] if len(r) > col]
tot = sum(float(r[col] or 0) for r in body) or 1.0
print(f"{len(body)} instructions, {int(tot)} samples")
order = {id(r): i for i, r in enumerate(body)}
for r in sorted(body, key=lambda r: -float(r[col] or 0))[:top_n]:
    print(f"{order[id(r)]:5d} {float(r[col] or 0):8.0f} {100 * float(r[col] or 0) / tot:5.1f}%  {r[1].strip()[:110]}")
