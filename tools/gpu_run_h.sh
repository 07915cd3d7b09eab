#!/bin/bash
# A/B of accumulator tile widths of the pair kernel (128 = four TMEM buffers, 256 = two), config 4 and a config-5 slice
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for t in 128 256; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity --param 8=$t > gpurun_out/h_bench_c4_t$t.json 2> gpurun_out/h_bench_c4_t$t.err
  timeout 600 python bench.py --config c5 --rows 125000000 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-parity --param 8=$t > gpurun_out/h_bench_c5slice_t$t.json 2> gpurun_out/h_bench_c5slice_t$t.err
done
for f in gpurun_out/h_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(d["ms_per_step"], r["frac"], r["step_frac"], r["phase_ms"], d["clocks"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done
