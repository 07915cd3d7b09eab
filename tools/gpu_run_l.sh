#!/bin/bash
# small-D path as two kernels (stream + fold): tests, configs 1/2 bench lines, launch list of config 1 and 2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -k "small or drop_in or golden or empty or last_labels or recolour or labels_to_host or out_of_core or kmeanspp or tie or nan" > gpurun_out/l_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/l_pytest.log
tail -4 gpurun_out/l_pytest.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(sys.argv[1], round(d["ms_per_step"],4), "%.3e"%d["value"], round(r["frac"],3), round(r["step_frac"],3), r["phase_ms"], d["config"]["checks"].get("parity",{}).get("ok"), d["gpu_launches"], d["clocks"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
for c in c1 c2; do
  timeout 300 python bench.py --config $c --steps 20 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/l_bench_$c.json 2> gpurun_out/l_bench_$c.err
  line gpurun_out/l_bench_$c.json
  ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/l_launches_$c.csv python bench.py --config $c --steps 6 --warmup 5 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/l_ncu_$c.log 2>&1
  grep -E "small_kernel|small_fold" gpurun_out/l_launches_$c.csv | tail -4 | awk -F'","' '{print $5, $NF}' | cut -c1-60,200-
done
