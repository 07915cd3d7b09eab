#!/usr/bin/env python
"""Launch list of the LAST Lloyd iteration in an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,
dram__bytes_write.sum --csv` log of bench.py: one line per launch of this library's kernels (time, DRAM bytes, grid)
and the iteration's totals.  The iteration is delimited by its first kernel (argument 2, a substring).

    python tools/ncu_iteration.py gpurun_out/n_launches_c3.csv prep_centroids_kernel
"""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
first = sys.argv[2]
launches = {}
order = []
for r in rows:
    lid = int(r[0])
    if lid not in launches:
        launches[lid] = {"name": r[4], "grid": r[8], "block": r[7]}
        order.append(lid)
    launches[lid][r[12]] = float(r[14].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "byte": 1e-9, "Kbyte": 1e-6,
                                                             "Mbyte": 1e-3, "Gbyte": 1.0}.get(r[13], 1.0)
ours = [i for i in order if "kmc::" in launches[i]["name"]] or list(order)      # (with ncu -k the names come without the namespace)
starts = [i for i in ours if first in launches[i]["name"]]
if not starts:
    raise SystemExit(f"no launch of {first}")
begin = starts[-2] if len(starts) > 1 else starts[-1]       # the last COMPLETE iteration
end = starts[-1] if len(starts) > 1 else ours[-1] + 1
it = [i for i in ours if begin <= i < end]
print(f"{'time':>10s} {'DRAM rd+wr':>14s}  {'grid':>14s} {'block':>12s}  kernel")
tt = tb = 0.0
for i in it:
    d = launches[i]
    t = d.get("gpu__time_duration.sum", 0.0)
    b = d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    tt += t
    tb += b
    name = d["name"].replace("void kmc::", "").split("(")[0]
    print(f"{t:8.4f} ms {b:11.4f} GB  {d['grid']:>14s} {d['block']:>12s}  {name}")
print(f"\niteration total: {tt:.4f} ms in {len(it)} launches, {tb:.4f} GB DRAM read+write")
