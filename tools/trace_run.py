"""Cycle trace of the tensor kernel's barrier hand-offs and the SM clock it really runs at.

Needs a library built with the trace compiled in:
    NVCC_EXTRA=-DKMC_TRACE python -c "from importlib import import_module as m; m('tensorflow-distributed-clustering_b200.build').build(force=True)"
    KMC_CHUNKS=1 python tools/trace_run.py
(rebuild without the flag afterwards).  Output: profiles/r1_tensor_kernel_bound.txt, sections 3 and 4.
"""
import ctypes
import os
import sys
from importlib import import_module

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
api = import_module("tensorflow-distributed-clustering_b200.api")
datagen = import_module("tensorflow-distributed-clustering_b200.datagen")

D, K = 128, 1024
for N in (4_000_000, 16_000_000, 100_000_000):
    X, _ = datagen.blobs_torch(N, D, K, 123, device="cuda:0")
    sh = api.LloydShard(N, D, K, "cuda:0")
    sh.bind_points(X)
    sh.set_centroids(X[:K].cpu().numpy())
    sh.iterate(3)
    sh.assign()
    torch.cuda.synchronize()
    t = np.zeros((256, 8), dtype=np.int64)
    assert sh.lib.kmc_debug_trace(ctypes.c_void_p(t.ctypes.data)) == 0
    dc, dt = t[255, 0] - t[0, 0], t[255, 1] - t[0, 1]
    print(f"N={N}: {dc / 255 * 4:.0f} cycles = {dt / 255 * 4:.0f} ns per 128-row tile -> SM clock in the kernel {dc / dt:.3f} GHz")
    if N == 16_000_000:
        t0 = t[40, 0]
        print("  step  wait_begin  t_full_seen  t_empty_arrive | row-tile end: start  done | issuer: buffer_ok  committed")
        for i in range(40, 52):
            r = t[i] - t0
            tail = f"{r[4]:8d} {r[5]:8d}" if i % 4 == 3 else " " * 17
            print(f"  {400 + i:4d}  {r[0]:10d}  {r[2]:11d}  {r[3]:14d} | {tail}          | {r[6]:10d} {r[7]:10d}")
    sh.close()
    del X, sh
    torch.cuda.empty_cache()
