#!/bin/bash
# 1-GPU validation of a build: full GPU test suite, smoke, config 4 and config 3 lines
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/d_pytest.log
tail -4 gpurun_out/d_pytest.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/d_bench_c4.json 2> gpurun_out/d_bench_c4.err
timeout 300 python bench.py --config c3 --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/d_bench_c3.json 2> gpurun_out/d_bench_c3.err
for f in gpurun_out/d_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(d["ms_per_step"], r["frac"], r["step_frac"], r["phase_ms"], d["clocks"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done
