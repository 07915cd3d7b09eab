#!/bin/bash
# ncu --set full: refine8 at config 4 (steady state), tensor kernel at config 3; launch lists of configs 1-3 with DRAM bytes
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD4="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity"
$CMD4 > gpurun_out/n_plain_c4.json 2> gpurun_out/n_plain_c4.err || echo "plain c4 failed"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:refine8_kernel --launch-skip 4 -c 1 -o gpurun_out/n_full_refine8_c4 -f $CMD4 > gpurun_out/n_ncu_refine8.log 2>&1; echo "ncu refine8 rc=$?"
CMD3="python bench.py --config c3 --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-parity"
$CMD3 > gpurun_out/n_plain_c3.json 2> gpurun_out/n_plain_c3.err || echo "plain c3 failed"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_assign_kernel --launch-skip 12 -c 1 -o gpurun_out/n_full_tc_c3 -f $CMD3 > gpurun_out/n_ncu_tc_c3.log 2>&1; echo "ncu tc c3 rc=$?"
for c in c1 c2 c3; do
  timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 300 --csv --log-file gpurun_out/n_launches_$c.csv python bench.py --config $c --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/n_ncu_l_$c.log 2>&1; echo "launch list $c rc=$?"
done
ls -la gpurun_out/n_*
