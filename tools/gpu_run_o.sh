#!/bin/bash
# incremental accumulation: changed rows through atomics (default) against the sort + segmented sum (--param 9=0)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/o_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/o_pytest.log
tail -3 gpurun_out/o_pytest.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(sys.argv[1], round(d["ms_per_step"],4), round(r["step_frac"],3), r["phase_ms"], d["config"]["checks"].get("parity",{}).get("ok"), d["config"]["checks"]["changed_rows_last_step"], d["gpu_launches"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
for c in c3 c4; do for v in default 0 default 0; do
  extra=""; [ $v = 0 ] && extra="--param 9=0"
  steps=10; [ $c = c3 ] && steps=20
  timeout 600 python bench.py --config $c --steps $steps --warmup 5 --no-e2e --no-cpu-baseline $extra > gpurun_out/o_bench_${c}_$v.json 2> gpurun_out/o_bench_${c}_$v.err; line gpurun_out/o_bench_${c}_$v.json
done; done
