#!/usr/bin/env python
"""Turn ncu exports into the short text summaries committed under profiles/.

    python tools/ncu_summary.py launches gpurun_out/launches_r1.csv
    python tools/ncu_summary.py full gpurun_out/r1_kernels_full.ncu-rep
"""
import collections
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size",
    "sm__cycles_elapsed.avg.per_second", "smsp__inst_executed.sum",
]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    data = [dict(zip(hdr, r)) for r in rows if len(r) == len(hdr) and r is not hdr and r[0] != "ID"]
    agg = collections.OrderedDict()
    for d in data:
        agg.setdefault(d["Kernel Name"].split("(")[0][:70], []).append(float(d["Metric Value"].replace(",", "")))
    total = sum(sum(v) for v in agg.values())
    print(f"{'kernel':72s} {'launches':>8s} {'avg ms':>10s} {'share':>7s}")
    for name, v in agg.items():
        print(f"{name:72s} {len(v):8d} {sum(v) / len(v) / 1e6:10.3f} {100 * sum(v) / total:6.1f}%")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        rec = dict(zip(hdr, r))
        print("=" * 100)
        print(rec.get("Kernel Name", "?")[:100])
        for k in KEYS:
            if k in rec:
                print(f"  {k:86s} {rec[k]:>16s} {units[hdr.index(k)]}")
        rd, wr = rec.get("dram__bytes_read.sum"), rec.get("dram__bytes_write.sum")
        if rd and wr:
            ur, uw = units[hdr.index("dram__bytes_read.sum")], units[hdr.index("dram__bytes_write.sum")]
            print(f"  dram traffic (read + write)  {rd} {ur} + {wr} {uw}")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
