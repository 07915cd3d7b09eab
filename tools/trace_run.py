import os, sys, ctypes
from importlib import import_module
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
api = import_module("tensorflow-distributed-clustering_b200.api")
datagen = import_module("tensorflow-distributed-clustering_b200.datagen")
N, D, K = 8_000_000, 128, 1024
X, _ = datagen.blobs_torch(N, D, K, 123, device="cuda:0")
sh = api.LloydShard(N, D, K, "cuda:0"); sh.bind_points(X); sh.set_centroids(X[:K].cpu().numpy())
sh.iterate(3)
lab = sh.assign(); torch.cuda.synchronize()
out = np.zeros((256, 16), dtype=np.int64)
rc = sh.lib.kmc_debug_trace(ctypes.c_void_p(out.ctypes.data)); assert rc == 0
t0 = out[40, 9]
print("warp 4 (half 0, quarter 0), cycles relative to step 440's wait begin")
print("step  wait_begin  t_full_ok    arrive | tail: start   end")
for i in range(40, 60):
    r = out[i]
    tail = "  ".join("%8d" % (r[k] - t0) for k in (10, 15)) if (i % 4) == 3 else ""
    print("%4d  %9d  %9d  %9d | %s" % (400 + i, r[9] - t0, r[2] - t0, r[4] - t0, tail))
print("cycles per row tile: %.0f" % ((out[200, 9] - out[40, 9]) / 40.0))
