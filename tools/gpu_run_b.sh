#!/bin/bash
# round-2 GPU batch B: CTA-pair tensor kernel (cta_group::2) -- correctness, then A/B timing at config 4
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or tensor or random_shapes or incremental or chunk or size_independent or streamed or refine or certified" > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -15 gpurun_out/b_pytest.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/b_bench_c4_pair.json 2> gpurun_out/b_bench_c4_pair.err
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity --param 7=0 > gpurun_out/b_bench_c4_nopair.json 2> gpurun_out/b_bench_c4_nopair.err
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/b_bench_c4_pair2.json 2> gpurun_out/b_bench_c4_pair2.err
timeout 300 python bench.py --config c3 --steps 20 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/b_bench_c3_pair.json 2> gpurun_out/b_bench_c3_pair.err
timeout 300 python bench.py --config c3 --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-parity --param 7=0 > gpurun_out/b_bench_c3_nopair.json 2> gpurun_out/b_bench_c3_nopair.err
for f in gpurun_out/b_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(d["ms_per_step"], d["value"], r["frac"], r["step_frac"], r["phase_ms"], d["config"]["checks"].get("parity"), d["clocks"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done
