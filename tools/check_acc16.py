#!/usr/bin/env python
"""fp16-accumulator tier of the tensor kernel (KMC_PARAM_TC_TERMS = 16) against float64: raw scores (layout of the
packed TMEM loads, rounding of the fp16 accumulation against the bound the kernel certifies with) and a short Lloyd
run against the oracle.  Needs a B200."""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

pkg = importlib.import_module("tensorflow-distributed-clustering_b200")
L = pkg._lib
dev = torch.device("cuda", 0)
ok_all = True
for (n, d, k) in [(4096, 128, 1024), (8192, 64, 4096), (5000, 32, 256), (3000, 96, 300), (2048, 256, 512)]:
    X, _ = pkg.datagen.blobs_numpy(n, d, max(2, k // 2), seed=7)
    C0 = X[:k].copy()
    shard = pkg.LloydShard(n, d, k, dev, path="tc")
    shard.bind_points(torch.from_numpy(X).to(dev))
    shard.set_centroids(C0)
    shard.set_param(L.KMC_PARAM_TC_TERMS, 16)
    S = shard.debug_tc_scores().double().cpu().numpy()                       # |x - c|^2, as the tier sees it
    X64, C64 = X.astype(np.float64), C0.astype(np.float64)
    T = (X64 * X64).sum(1)[:, None] + (C64 * C64).sum(1)[None, :] - 2.0 * X64 @ C64.T       # the tier adds |x|^2 through the MMA
    err = np.abs(S - T)
    xn = np.sqrt((X64 * X64).sum(1))
    cn = np.sqrt((C64 * C64).sum(1))
    steps = (-(-d // 64) * 64) // 16
    bound = 1.25 * steps / 2048.0 * 2.0 * xn[:, None] * cn.max() + 1.25 / 2048.0 * np.abs(T)     # the accumulation part of the kernel's bound
    ratio = (err / bound).max()
    print(f"n={n} D={d} K={k}: max |score error| {err.max():.4g} (scores up to {np.abs(T).max():.4g}), "
          f"max error / accumulation bound {ratio:.3f}, mean {np.mean(err / bound):.4f}")
    ok_all &= bool(ratio < 0.8)
    # a short run through the product loop with the tier forced
    n_iter, inertia, _ = shard.iterate(3)
    C = shard.get_centroids()
    labels = shard.assign().cpu().numpy()
    st = shard.stats()
    shard.close()
    ref = oracle.run_lloyd(X, C0, 3, form="gemm")
    bad = oracle.label_mismatches(labels, ref["cluster_idx"], ref["final_d1"], ref["final_d2"])
    errc = oracle.rel_err(C, ref["end_center"])
    erri = abs(inertia - ref["history"][-1]["inertia"]) / ref["history"][-1]["inertia"]
    print(f"    3 iterations: centroid rel err {errc:.2e}, inertia rel err {erri:.2e}, label mismatches {len(bad)}, "
          f"refined rows {st['ambiguous']} of {n}")
    ok_all &= bool(errc <= 1e-4 and erri <= 1e-4 and len(bad) == 0)
print("ACC16 OK" if ok_all else "ACC16 FAILED")
