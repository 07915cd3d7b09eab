#!/bin/bash
# A/B of library variants built into tools/var_*.so: swaps each in, runs the C4 bench, restores.
# CHS="1 4" row-chunk settings to run; MODES="aug plain"
L=tensorflow-distributed-clustering_b200/libkmc_b200.so
cp $L /tmp/base.so
run() { for mode in ${MODES:-aug}; do for ch in ${CHS:-1}; do
  if [ $mode = plain ]; then export KMC_NO_BIAS_MMA=1; else unset KMC_NO_BIAS_MMA; fi
  echo -n "$1 $mode chunks=$ch: "; KMC_CHUNKS=$ch timeout 200 python bench.py --steps ${STEPS:-6} --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python tools/benchline.py; done; done; }
run base
for v in tools/var_*.so; do cp $v $L; run $(basename $v .so); done
cp /tmp/base.so $L
