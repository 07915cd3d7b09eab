#!/bin/bash
# round-2 GPU batch A: tests, smoke, first bench lines with the incremental accumulation
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/a_gpu.txt 2>&1
free -g >> gpurun_out/a_gpu.txt; nproc >> gpurun_out/a_gpu.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/a_smoke.log 2>&1; tail -4 gpurun_out/a_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench_c4.json 2> gpurun_out/a_bench_c4.err; tail -c 3000 gpurun_out/a_bench_c4.json
timeout 300 python bench.py --steps 10 --warmup 3 --param 5=4 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/a_bench_c4_chunks4.json 2> gpurun_out/a_bench_c4_chunks4.err
timeout 300 python bench.py --steps 10 --warmup 3 --param 6=0 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/a_bench_c4_noinc.json 2> gpurun_out/a_bench_c4_noinc.err
for c in c3 c2 c1; do
  timeout 300 python bench.py --config $c --steps 20 --warmup 5 > gpurun_out/a_bench_$c.json 2> gpurun_out/a_bench_$c.err
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/a_launches_c4.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/a_ncu.log 2>&1
for f in gpurun_out/a_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(d["ms_per_step"], d["value"], r["frac"], r["step_frac"], r["phase_ms"], d["config"]["checks"], d.get("e2e",{}) and d["e2e"].get("value"), d.get("e2e_pageable_f64",{}) and d["e2e_pageable_f64"].get("value"))
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done
