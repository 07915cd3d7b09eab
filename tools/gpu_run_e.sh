#!/bin/bash
# multi-GPU batch: $1 = number of GPUs; multi-GPU tests (N >= 2), then bench lines of the configs given as further arguments
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
G=$1; shift
if [ "$G" = "2" ]; then
  timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/e_pytest_multi_${G}gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/e_pytest_multi_${G}gpu.log
  tail -4 gpurun_out/e_pytest_multi_${G}gpu.log
fi
PORT=29500
for c in "$@"; do
  PORT=$((PORT+1))
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port $PORT \
     bench.py --gpus $G --config $c --steps 10 --warmup 3 > gpurun_out/e_bench_${c}_${G}gpu.json 2> gpurun_out/e_bench_${c}_${G}gpu.err
  grep -v "^\[W\|^W\|^\*\*\*" gpurun_out/e_bench_${c}_${G}gpu.json | tail -1 > gpurun_out/e_tmp && python - gpurun_out/e_tmp <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; print(d["config"]["workload"][:3], d["n_gpus"], d["ms_per_step"], d["value"], r["frac"], r["step_frac"], r["phase_ms"], d["config"]["checks"], d["e2e"] and d["e2e"]["value"], d["clocks"])
except Exception as e:
    print("ERR", e)
PY
  tail -c 1500 gpurun_out/e_bench_${c}_${G}gpu.err | grep -v Warning | tail -8
done
