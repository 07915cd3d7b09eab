#!/bin/bash
# speed envelope of config 4 on harder data (not the headline), a one-GPU slice of config 5, launch list of config 4
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; c=d["config"]; print(sys.argv[1], c.get("data_variant"), round(d["ms_per_step"],3), round(r["frac"],3), round(r["step_frac"],3), r["phase_ms"], "refined", c["refined_rows_last_step"], "overflow", c["walk_overflow_rows_last_step"], c["checks"].get("parity",{}).get("ok"), d["clocks"]["sm_mhz"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
for v in "--data blobs4" "--data uniform" "--sigma 4" "--sigma 2"; do
  tag=$(echo $v | tr -d ' -' )
  timeout 900 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline $v > gpurun_out/p_bench_c4_$tag.json 2> gpurun_out/p_bench_c4_$tag.err
  line gpurun_out/p_bench_c4_$tag.json
done
timeout 900 python bench.py --config c5 --rows 125000000 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/p_bench_c5slice.json 2> gpurun_out/p_bench_c5slice.err; line gpurun_out/p_bench_c5slice.json
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/p_launches_c4.csv $CMD > gpurun_out/p_ncu_c4.log 2>&1; echo "launch list c4 rc=$?"
