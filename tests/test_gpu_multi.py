"""Multi-GPU tests (need >= 2 CUDA devices; skipped otherwise): the single-process form of the
reference's call -- `GPU_names` with several devices (reference :180-263) -- against the oracle."""
import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,d,k,g", [(30001, 2, 8, 2), (20000, 64, 128, 2)])
def test_single_process_multi_gpu_matches_oracle(n, d, k, g, pkg, cuda_device):
    import torch

    if torch.cuda.device_count() < g:
        pytest.skip(f"needs {g} GPUs")
    X, _ = pkg.datagen.blobs_numpy(n, d, k, seed=123128)
    C0 = X[:k].copy()
    names = [f"/device:GPU:{i}" for i in range(g)]
    res = pkg.distribuited_k_means(X, k, names, C0, 6)
    ref = oracle.run_lloyd(X, C0, 6, n_shards=g, form="gemm" if d * k > 4096 else "diff")
    assert oracle.rel_err(res["end_center"], ref["end_center"]) <= 1e-4
    assert len(oracle.label_mismatches(res["cluster_idx"], ref["cluster_idx"], ref["final_d1"], ref["final_d2"])) == 0
    last = pkg.shard_sizes(n, g)[-1]
    # `centroids`: labels of the last GPU's shard in the final iteration, before its update (:279, :286)
    h = ref["history"][-1]
    assert res["centroids"].shape == (last,)
    assert len(oracle.label_mismatches(res["centroids"], ref["centroids"], h["d1"][-last:], h["d2"][-last:])) == 0
    assert abs(res["inertia"] - ref["history"][-1]["inertia"]) <= 1e-4 * ref["history"][-1]["inertia"]


@pytest.mark.parametrize("n,d,k", [(40001, 64, 128), (50000, 3, 16)])
def test_torchrun_nccl_spmd_matches_oracle(n, d, k, pkg, cuda_device, tmp_path):
    """One process per GPU under torchrun, world size 2, NCCL all-reduce of the partial buffers
    (`distribuited_k_means_spmd`, the form bench.py --gpus N times): the gathered result against the oracle run with
    two shards -- the reference's G-invariance (:184, :257-258) -- and replicated centroids bit-identical."""
    import os
    import socket
    import subprocess
    import sys

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = tmp_path / "spmd.npz"
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_spmd_worker.py")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), worker, str(out), str(n), str(d), str(k), "5"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    z = np.load(out)
    X, _ = pkg.datagen.blobs_numpy(n, d, k, seed=123128)
    ref = oracle.run_lloyd(X, X[:k].copy(), 5, n_shards=2, form="gemm" if d * k > 4096 else "diff")
    assert bool(z["end_centers_equal"])
    assert oracle.rel_err(z["end_center"], ref["end_center"]) <= 1e-4
    assert len(oracle.label_mismatches(z["cluster_idx"], ref["cluster_idx"], ref["final_d1"], ref["final_d2"])) == 0
    last = pkg.shard_sizes(n, 2)[-1]
    h = ref["history"][-1]
    assert len(oracle.label_mismatches(z["last_centroids"], ref["centroids"], h["d1"][-last:], h["d2"][-last:])) == 0
    assert abs(float(z["inertia"]) - h["inertia"]) <= 1e-4 * h["inertia"]
