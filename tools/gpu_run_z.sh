#!/bin/bash
# ncu --set full of one steady-state iteration's tensor-kernel launches at config 4 (the fp16-accumulator pair kernel and its gated partners)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 6 --no-e2e --no-cpu-baseline --no-parity"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_assign_kernel --launch-skip 28 -c 4 -o gpurun_out/z_full_tc_c4 -f $CMD > gpurun_out/z_ncu.log 2>&1; echo "ncu rc=$?"
python -m pytest tests -m gpu -x -q -k "fp16 or tensor_path_with or well_separated" 2>&1 | tail -2
