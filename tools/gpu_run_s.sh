#!/bin/bash
# ncu --set full of the fp16-accumulator tensor kernel at config 4 (source-level stalls)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity --param 3=16"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_assign_kernel --launch-skip 4 -c 1 -o gpurun_out/s_full_tc16_c4 -f $CMD > gpurun_out/s_ncu.log 2>&1; echo "ncu rc=$?"
