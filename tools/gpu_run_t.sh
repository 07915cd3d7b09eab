#!/bin/bash
# fp16-accumulator tier as the adaptive policy's first tier: all GPU tests, smoke, configs 3/4 + c5 slice, envelope
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_pytest.log
tail -6 gpurun_out/t_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/t_smoke.log 2>&1; tail -2 gpurun_out/t_smoke.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; c=d["config"]; print(sys.argv[1], c.get("data_variant",{}).get("data"), c.get("data_variant",{}).get("sigma"), round(d["ms_per_step"],3), round(r["frac"],3), round(r["step_frac"],3), r["phase_ms"], "refined", c["refined_rows_last_step"], "overflow", c["walk_overflow_rows_last_step"], c["checks"].get("parity",{}).get("ok"), d["clocks"]["sm_mhz"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
timeout 900 python bench.py --no-e2e --no-cpu-baseline > gpurun_out/t_bench_c4.json 2> gpurun_out/t_bench_c4.err; line gpurun_out/t_bench_c4.json
timeout 900 python bench.py --config c5 --rows 125000000 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/t_bench_c5slice.json 2> gpurun_out/t_bench_c5slice.err; line gpurun_out/t_bench_c5slice.json
for v in "--data uniform"; do
  tag=$(echo $v | tr -d ' -' )
  timeout 900 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline $v > gpurun_out/t_bench_c4_$tag.json 2> gpurun_out/t_bench_c4_$tag.err
  line gpurun_out/t_bench_c4_$tag.json
done
