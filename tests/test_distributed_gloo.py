"""world_size-2 `gloo` test (CPU) of the multi-process host logic: each rank owns the rows the
reference's array_split rule gives it (reference :184), runs the shared `lloyd_loop` on its own
shard, and the partial buffers are summed with torch.distributed.all_reduce between accumulate and
update.  The CUDA kernels cannot run here, so the shard engine is a test double that fills the same
partial-buffer layout ([K*D sums | K counts | inertia | reserved], float64) from the oracle; what is
under test is the product's loop, sharding and reduction protocol, not the arithmetic."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleShard:
    """Test double with the LloydShard interface used by lloyd_loop."""

    def __init__(self, X, C0):
        import oracle
        self.oracle = oracle
        self.X = X
        self.C = np.asarray(C0, dtype=np.float64).copy()
        K, D = self.C.shape
        self.K, self.D = K, D
        self.partials = torch.zeros(K * D + K + 2, dtype=torch.float64)
        self.shift2 = float("nan")

    def accumulate(self):
        it = self.oracle.lloyd_iteration(self.X, self.C) if len(self.X) else None
        self.partials.zero_()
        if it is not None:
            K, D = self.K, self.D
            self.partials[:K * D] = torch.from_numpy(it["sums"].reshape(-1))
            self.partials[K * D:K * D + K] = torch.from_numpy(it["counts"])
            self.partials[K * D + K] = it["inertia"]

    def update(self):
        K, D = self.K, self.D
        p = self.partials.numpy()
        sums, counts = p[:K * D].reshape(K, D), p[K * D:K * D + K]
        new = np.where(counts[:, None] > 0, sums / np.maximum(counts[:, None], 1), self.C)
        self.shift2 = float(((new - self.C) ** 2).sum())
        self.C = new

    def stats(self):
        return {"inertia": float(self.partials[self.K * self.D + self.K]), "shift2": self.shift2}


def _worker(rank, world, port, n, d, k, iters, tol, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import importlib
    pkg = importlib.import_module("tensorflow-distributed-clustering_b200")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    X, _ = pkg.datagen.blobs_numpy(n, d, k, seed=42)
    r0, r1 = pkg.shard_ranges(n, world)[rank]
    shard = OracleShard(X[r0:r1], X[:k])
    n_iter, inertia, shift2 = pkg.lloyd_loop([shard], iters, tol,
                                             reduce_fn=lambda s: dist.all_reduce(s[0].partials))
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), C=shard.C, n_iter=n_iter, inertia=inertia, shift2=shift2)
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("n,tol", [(1001, 0.0), (1500, 1e-9)])
def test_two_rank_gloo_loop_matches_single_process_oracle(n, tol, tmp_path, pkg):
    import oracle

    d, k, iters = 3, 5, (12 if tol == 0 else 100)
    mp.spawn(_worker, args=(2, _free_port(), n, d, k, iters, tol, str(tmp_path)), nprocs=2, join=True)
    r = [np.load(tmp_path / f"rank{i}.npz") for i in range(2)]
    assert np.array_equal(r[0]["C"], r[1]["C"]), "ranks must hold identical centroids without a broadcast"
    assert int(r[0]["n_iter"]) == int(r[1]["n_iter"])
    X, _ = pkg.datagen.blobs_numpy(n, d, k, seed=42)
    ref = oracle.run_lloyd(X, X[:k], iters, n_shards=2, tol=tol)
    assert int(r[0]["n_iter"]) == ref["n_iter"]
    assert oracle.rel_err(r[0]["C"], ref["end_center"]) < 1e-12
    assert abs(float(r[0]["inertia"]) - ref["history"][-1]["inertia"]) < 1e-9 * ref["history"][-1]["inertia"]
    if tol > 0:
        assert ref["n_iter"] < iters
