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
    assert np.array_equal(res["centroids"], res["cluster_idx"][-last:])      # labels of the last GPU's shard (:279)
    assert abs(res["inertia"] - ref["history"][-1]["inertia"]) <= 1e-4 * ref["history"][-1]["inertia"]
