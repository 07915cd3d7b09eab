"""Where the end-to-end time of distribuited_k_means goes at config 4 (one GPU): wall clock per stage."""
import os, sys, time
from importlib import import_module
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
api = import_module("tensorflow-distributed-clustering_b200.api")
datagen = import_module("tensorflow-distributed-clustering_b200.datagen")

N, D, K, IT = int(os.environ.get("ROWS", 100_000_000)), 128, 1024, 10
dev = torch.device("cuda:0")
X, _ = datagen.blobs_torch(N, D, K, 123, device=dev)
host = torch.empty((N, D), dtype=torch.float32, pin_memory=True)
host.copy_(X); torch.cuda.synchronize()
C0 = X[:K].cpu().numpy()
del X; torch.cuda.empty_cache()

def tick(label, t0):
    torch.cuda.synchronize(); t = time.perf_counter(); print("%-28s %8.1f ms" % (label, (t - t0) * 1e3)); return t

t = time.perf_counter(); t_all = t
sh = api.LloydShard(N, D, K, dev); t = tick("LloydShard (workspace)", t)
sh.stage_from_host(host, 0, None); t = tick("stage_from_host", t)
sh.set_centroids(C0); t = tick("set_centroids", t)
sh.iterate(1); t = tick("iteration 1", t)
sh.iterate(1); t = tick("iteration 2", t)
sh.iterate(IT - 2); t = tick("iterations 3..%d" % IT, t)
c = sh.get_centroids(); t = tick("get_centroids", t)
lab = sh.assign(); t = tick("assign", t)
res = np.empty(N, dtype=np.int64); sh.labels_to_host(lab, res); t = tick("labels_to_host (int64)", t)
assert (torch.from_numpy(res[:1000000]).to(dev) == lab[:1000000]).all()
sh.close(); t = tick("close", t)
print("total %.1f ms" % ((t - t_all) * 1e3))
del lab, sh; torch.cuda.empty_cache()
def wrap(name):
    f = getattr(api.LloydShard, name)
    def g(self, *a, **k):
        torch.cuda.synchronize(); t0 = time.perf_counter(); r = f(self, *a, **k); torch.cuda.synchronize()
        print("   [%s] %.1f ms" % (name, (time.perf_counter() - t0) * 1e3)); return r
    setattr(api.LloydShard, name, g)
for nm in ("assign", "labels_to_host", "close", "get_centroids", "iterate", "stage_from_host", "__init__"):
    wrap(nm)
t = time.perf_counter()
r = api.distribuited_k_means(host, K, ["/device:GPU:0"], C0, IT)
print("distribuited_k_means total %.1f ms (setup %.1f, init %.1f, compute %.1f)" % ((time.perf_counter() - t) * 1e3,
      r["setup_time"] * 1e3, r["initialization_time"] * 1e3, r["computation_time"] * 1e3))
