#!/bin/bash
# speed envelope on harder data + ncu captures (launch list, whole-iteration DRAM bytes, full set of the tensor kernel)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for v in "--data blobs4" "--data uniform" "--sigma 4" "--sigma 1"; do
  tag=$(echo $v | tr -d ' -' )
  timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline $v > gpurun_out/g_bench_c4_$tag.json 2> gpurun_out/g_bench_c4_$tag.err
  python - gpurun_out/g_bench_c4_$tag.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; c=d["config"]; print(c.get("data_variant"), d["ms_per_step"], r["step_frac"], r["phase_ms"], c["refined_rows_last_step"], c["checks"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done
