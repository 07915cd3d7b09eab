#!/bin/bash
# small-D path: tests and config 1 / 2 lines
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "small or drop_in or golden or empty or last_labels or recolour or labels_to_host or out_of_core or kmeanspp" > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest.log
tail -6 gpurun_out/f_pytest.log
for c in c1 c2; do
  timeout 300 python bench.py --config $c --steps 20 --warmup 5 > gpurun_out/f_bench_$c.json 2> gpurun_out/f_bench_$c.err
  python - gpurun_out/f_bench_$c.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(d["config"]["workload"][:3], d["ms_per_step"], d["value"], r["frac"], r["step_frac"], r["phase_ms"], d["config"]["checks"].get("parity"), d["e2e"] and d["e2e"]["value"], d["gpu_launches"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done
