#!/bin/bash
# confirm run of the committed state: all GPU tests, smoke, the driver's default bench line, configs 1-3
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/i_pytest.log
tail -4 gpurun_out/i_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/i_smoke.log 2>&1; tail -4 gpurun_out/i_smoke.log
timeout 900 python bench.py > gpurun_out/i_bench_c4.json 2> gpurun_out/i_bench_c4.err; echo "bench rc=$?"
for c in c1 c2 c3; do
  timeout 300 python bench.py --config $c --steps 20 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/i_bench_$c.json 2> gpurun_out/i_bench_$c.err
done
for f in gpurun_out/i_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(d["ms_per_step"], d["value"], r["frac"], r["step_frac"], r["phase_ms"], d["config"]["checks"].get("parity",{}).get("ok"), d["e2e"] and d["e2e"]["value"], d["gpu_launches"], d["clocks"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done
