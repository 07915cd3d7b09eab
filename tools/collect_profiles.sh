#!/bin/bash
# Round-1 evidence: bench lines, ncu launch list and ncu --set full summaries (run on the GPU box).
# Every ncu pass repeats a command that has just exited 0 without ncu.
set -u
O=gpurun_out/prof; mkdir -p $O
python bench.py > $O/bench_c4_1gpu.json 2> $O/bench_c4_1gpu.err || exit 1
for c in c1 c2 c3 c5; do python bench.py --config $c --no-e2e --no-cpu-baseline > $O/bench_$c.json 2>/dev/null; done
export KMC_CHUNKS=1
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD > $O/plain_run.json 2>/dev/null || exit 2
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"tc_assign_kernel|refine|segsum" --launch-skip 24 -c 6 -o $O/full_c4 -f $CMD > $O/ncu_full.log 2>&1
echo done
