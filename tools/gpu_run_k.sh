#!/bin/bash
# ncu --set full of the small-D kernel at configs 1 and 2 (source-level stall sampling), after the plain run exited 0
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for c in c1 c2; do
  CMD="python bench.py --config $c --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-parity"
  $CMD > gpurun_out/k_plain_$c.json 2> gpurun_out/k_plain_$c.err || { echo "plain run failed $c"; continue; }
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:small_kernel --launch-skip 4 -c 1 -o gpurun_out/k_full_$c -f $CMD > gpurun_out/k_ncu_$c.log 2>&1
  echo "ncu $c rc=$?"
done
ls -la gpurun_out/*.ncu-rep
