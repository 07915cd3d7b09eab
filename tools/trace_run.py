import os, sys, ctypes
from importlib import import_module
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
api = import_module("tensorflow-distributed-clustering_b200.api")
datagen = import_module("tensorflow-distributed-clustering_b200.datagen")
for N in (4_000_000, 16_000_000, 100_000_000):
    D, K = 128, 1024
    X, _ = datagen.blobs_torch(N, D, K, 123, device="cuda:0")
    sh = api.LloydShard(N, D, K, "cuda:0"); sh.bind_points(X); sh.set_centroids(X[:K].cpu().numpy())
    sh.iterate(3)
    sh.assign(); torch.cuda.synchronize()
    out = np.zeros((256, 16), dtype=np.int64)
    rc = sh.lib.kmc_debug_trace(ctypes.c_void_p(out.ctypes.data)); assert rc == 0
    last = 255 if N > 20_000_000 else min(255, (N // 128 // 148) * 4 - 401 - 8)
    dc, dt = out[last, 0] - out[0, 0], out[last, 1] - out[0, 1]
    print("N=%d: steps 400..%d: %d cycles in %d ns -> SM clock %.3f GHz; %.0f cycles = %.0f ns per row tile" %
          (N, 400 + last, dc, dt, dc / dt, dc / last * 4, dt / last * 4))
    sh.close(); del X, sh; torch.cuda.empty_cache()
