#!/bin/bash
# final 1-GPU evidence of the round: bench lines of configs 1-4 (with e2e legs, parity, CPU arms), reference arm,
# steady-state launch list of config 4, ncu --set full of the fp16-accumulator pair kernel
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(sys.argv[1], round(d["ms_per_step"],4), "%.3e"%d["value"], round(r["frac"],3), round(r["step_frac"],3), r["phase_ms"], d["config"]["checks"].get("parity",{}).get("ok"), d["e2e"] and "%.3e"%d["e2e"]["value"], d["gpu_launches"], d["clocks"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
timeout 900 python bench.py > gpurun_out/u_bench_c4.json 2> gpurun_out/u_bench_c4.err; line gpurun_out/u_bench_c4.json
for c in c1 c2 c3; do
  timeout 600 python bench.py --config $c --steps 20 --warmup 5 > gpurun_out/u_bench_$c.json 2> gpurun_out/u_bench_$c.err; line gpurun_out/u_bench_$c.json
done
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/u_bench_c4_reference.json 2> gpurun_out/u_bench_c4_reference.err
K="tc_assign|refine|simt_assign|hist_kernel|scan_kernel|scatter|segsum|diff_labels|delta|update_kernel|prep_centroids|ccdist|nbrsort|compact_worklist|tc_gate"
CMD="python bench.py --steps 3 --warmup 6 --no-e2e --no-cpu-baseline --no-parity"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"$K" -c 400 --csv --log-file gpurun_out/u_launches_c4.csv $CMD > gpurun_out/u_ncu_l.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_assign_kernel --launch-skip 21 -c 3 -o gpurun_out/u_full_tc16_c4 -f $CMD > gpurun_out/u_ncu_f.log 2>&1; echo "ncu full rc=$?"
