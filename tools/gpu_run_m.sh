#!/bin/bash
# full confirm of the working tree: all GPU tests, smoke, configs 1-3 (with e2e legs) and the driver's default line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/m_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/m_pytest.log
tail -4 gpurun_out/m_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/m_smoke.log 2>&1; tail -4 gpurun_out/m_smoke.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(sys.argv[1], round(d["ms_per_step"],4), "%.3e"%d["value"], round(r["frac"],3), round(r["step_frac"],3), r["phase_ms"], d["config"]["checks"].get("parity",{}).get("ok"), d["e2e"] and "%.3e"%d["e2e"]["value"], d["gpu_launches"], d["clocks"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
for c in c1 c2 c3; do
  timeout 600 python bench.py --config $c --steps 20 --warmup 5 > gpurun_out/m_bench_$c.json 2> gpurun_out/m_bench_$c.err; line gpurun_out/m_bench_$c.json
done
timeout 900 python bench.py > gpurun_out/m_bench_c4.json 2> gpurun_out/m_bench_c4.err; line gpurun_out/m_bench_c4.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/m_bench_c4_reference.json 2> gpurun_out/m_bench_c4_reference.err; tail -c 600 gpurun_out/m_bench_c4_reference.json
