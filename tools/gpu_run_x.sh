#!/bin/bash
# last confirm of the committed tree: all GPU tests, smoke, the driver's default bench line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/x_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/x_pytest.log
tail -4 gpurun_out/x_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/x_smoke.log 2>&1; tail -4 gpurun_out/x_smoke.log
timeout 900 python bench.py > gpurun_out/x_bench_c4.json 2> gpurun_out/x_bench_c4.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/x_bench_c4.json').read().strip().splitlines()[-1])
r=d["roofline"]; print(round(d["ms_per_step"],3), "%.3e"%d["value"], round(r["frac"],3), round(r["step_frac"],3), r["phase_ms"], d["config"]["checks"]["parity"]["ok"], "%.3e"%d["e2e"]["value"], d["gpu_launches"], d["clocks"])
PY
