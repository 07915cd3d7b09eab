#!/bin/bash
# A/B of library builds under build_variants/*.so in the timed loop of bench.py (power-capped regime), alternating;
# prints the phase times.  Arguments: passed on to bench.py.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=tensorflow-distributed-clustering_b200/libkmc_b200.so
cp $L /tmp/keep.so
for rep in 1 2; do for v in build_variants/*.so; do
  cp $v $L; tag=$(basename $v .so)
  timeout 600 python bench.py "$@" --no-e2e --no-cpu-baseline --no-parity > gpurun_out/abb_$tag.json 2> gpurun_out/abb_$tag.err
  python - gpurun_out/abb_$tag.json $tag <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], round(d["ms_per_step"],3), d["roofline"]["phase_ms"], d["clocks"]["sm_mhz"])
except Exception as e:
    print(sys.argv[2], "ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
done; done
cp /tmp/keep.so $L
