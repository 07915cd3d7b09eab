#!/bin/bash
# A/B of builds of the CTA-pair tensor kernel at config 4 (wait-hint / spin variants)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
P=tensorflow-distributed-clustering_b200
cp $P/libkmc_b200.so /tmp/lib_main.so
for v in "$@"; do
  cp build_variants/lib_$v.so $P/libkmc_b200.so
  timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/c_bench_c4_$v.json 2> gpurun_out/c_bench_c4_$v.err
done
cp /tmp/lib_main.so $P/libkmc_b200.so
for f in gpurun_out/c_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(d["ms_per_step"], r["frac"], r["step_frac"], r["phase_ms"], d["clocks"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done
