#!/bin/bash
# slot-table flush of the small-D kernel + pipelined refine8: tests, configs 1/2, A/B of the refine against the previous build
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=tensorflow-distributed-clustering_b200/libkmc_b200.so
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/j_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j_pytest.log
tail -4 gpurun_out/j_pytest.log
for c in c1 c2; do
  timeout 300 python bench.py --config $c --steps 20 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/j_bench_$c.json 2> gpurun_out/j_bench_$c.err
done
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(sys.argv[1], round(d["ms_per_step"],4), "%.3e"%d["value"], round(r["frac"],3), round(r["step_frac"],3), r["phase_ms"], d["config"]["checks"].get("parity",{}).get("ok"), d["gpu_launches"], d["clocks"])
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
line gpurun_out/j_bench_c1.json; line gpurun_out/j_bench_c2.json
cp $L /tmp/new.so
for v in new old new old; do
  if [ $v = old ]; then cp build_variants/var_oldrefine.so $L; else cp /tmp/new.so $L; fi
  timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/j_bench_c4_$v.json 2> gpurun_out/j_bench_c4_$v.err
  line gpurun_out/j_bench_c4_$v.json
done
cp /tmp/new.so $L
timeout 300 python bench.py --config c3 --steps 20 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/j_bench_c3.json 2> gpurun_out/j_bench_c3.err; line gpurun_out/j_bench_c3.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/j_launches_c1.csv python bench.py --config c1 --steps 10 --warmup 5 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/j_ncu_c1.log 2>&1
grep small_kernel gpurun_out/j_launches_c1.csv | tail -5
