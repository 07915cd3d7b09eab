#!/bin/bash
# A/B of library builds under build_variants/*.so on one kernel: ncu gpu__time_duration of the launches matching $1
# (regex) in `bench.py $2...`; per-launch times under ncu are serialised and unthrottled, so builds compare cleanly.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
K=$1; shift
L=tensorflow-distributed-clustering_b200/libkmc_b200.so
cp $L /tmp/keep.so
for v in build_variants/*.so; do
  cp $v $L
  tag=$(basename $v .so)
  timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum --clock-control none -k regex:$K -c 40 --csv --log-file gpurun_out/ab_$tag.csv python bench.py "$@" --no-e2e --no-cpu-baseline --no-parity > gpurun_out/ab_$tag.log 2>&1
  python - gpurun_out/ab_$tag.csv $tag <<'PY'
import csv,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10 and r[0].isdigit() and r[-3]=="gpu__time_duration.sum"]
print(sys.argv[2], [round(float(r[-1])/1e3,1) for r in rows if float(r[-1])>20000][:12])
PY
done
cp /tmp/keep.so $L
