#!/bin/bash
# A/B of library variants built into tools/var_*.so: swaps each in, runs the C4 bench (assign phase time), restores.
L=tensorflow-distributed-clustering_b200/libkmc_b200.so
cp $L /tmp/base.so
run() { for mode in aug plain; do
  if [ $mode = plain ]; then export KMC_NO_BIAS_MMA=1; else unset KMC_NO_BIAS_MMA; fi
  echo -n "$1 $mode: "; KMC_CHUNKS=${CH:-1} timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python tools/benchline.py; done; }
run base
for v in tools/var_*.so; do cp $v $L; run $(basename $v .so); done
cp /tmp/base.so $L
