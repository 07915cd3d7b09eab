"""GPU test of kmc_cluster_sums_noalloc, the one entry point of libkmc_b200 that allocates nothing.

Who needs it: hosts that must not call cudaMalloc -- an operator that a framework runs with only its
input and output buffers because the framework's pool owns all device memory, or a serving loop on a
fixed memory budget.  Such a caller cannot hand the library a workspace, so the scratch of the sorted
accumulation sits in a fixed-size __device__ array inside the library (limits: 1 048 576 rows and 4096
clusters per call).

The test calls the new entry point through the C ABI twice in a row, on two different problems, and
checks each result against what the existing, workspace-based entry point (kmc_accumulate) leaves in
its partial buffer for the same points, labels and centroids: counts exactly, sums and inertia within
the 1e-4 relative tolerance the other GPU tests use.  The second call runs on a smaller problem with a
different number of clusters, so anything left in the shared scratch by the first call would show up.
Further cases check it directly against NumPy (bincount, float64 per-cluster sums, inertia): at exactly
the largest sizes it accepts, with a row length that rules out 16-byte loads, and with a few labels
outside [0, K), whose rows must simply be left out.
This file is named so that pytest collects it last; it needs nothing from the other tests.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-4


def _reference_partials(pkg, X, C, device, path):
    """Existing entry points: labels from kmc_assign, partial buffer from kmc_accumulate."""
    import torch

    s = pkg.LloydShard(X.shape[0], X.shape[1], C.shape[0], device, path=path)
    pts = torch.from_numpy(X).to(device)
    s.bind_points(pts)
    s.set_centroids(C)
    labels = s.assign().clone()
    s.accumulate()
    torch.cuda.synchronize(device)
    part = s.partials.clone()
    cent = s.centroids_dev.clone()
    s.close()
    return pts, labels, cent, part


def _noalloc_partials(pkg, pts, labels, cent, device):
    import torch

    lib = pkg._lib.load()
    n, D = pts.shape
    K = cent.shape[0]
    out = torch.full((K * D + K + 2,), float("nan"), dtype=torch.float64, device=device)   # must be overwritten
    stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
    rc = lib.kmc_cluster_sums_noalloc(device.index or 0, ctypes.c_void_p(pts.data_ptr()),
                                      ctypes.c_void_p(labels.data_ptr()), ctypes.c_void_p(cent.data_ptr()),
                                      n, D, K, ctypes.c_void_p(out.data_ptr()), stream)
    assert rc == 0, lib.kmc_last_error()
    torch.cuda.synchronize(device)
    return out


def _compare(got, want, K, D):
    got, want = got.cpu().numpy(), want.cpu().numpy()
    sums_g, sums_w = got[:K * D], want[:K * D]
    assert np.array_equal(got[K * D:K * D + K], want[K * D:K * D + K]), "counts differ"
    assert np.max(np.abs(sums_g - sums_w)) <= RTOL * np.max(np.abs(sums_w))
    assert abs(got[K * D + K] - want[K * D + K]) <= RTOL * abs(want[K * D + K])


def test_noalloc_matches_workspace_path_on_two_calls_in_a_row(pkg, cuda_device):
    problems = [(200_000, 24, 300, "simt", 5), (37_013, 128, 64, "tc", 9)]
    prepared = []
    for n, d, k, path, seed in problems:
        X, _ = pkg.datagen.blobs_numpy(n, d, k, seed=seed)
        prepared.append(_reference_partials(pkg, X, X[:k].copy(), cuda_device, path) + (k, d))
    # two calls back to back on the shared scratch, second one smaller and with another K
    for pts, labels, cent, want, k, d in prepared:
        got = _noalloc_partials(pkg, pts, labels, cent, cuda_device)
        _compare(got, want, k, d)
    # and once more in the opposite order
    for pts, labels, cent, want, k, d in reversed(prepared):
        _compare(_noalloc_partials(pkg, pts, labels, cent, cuda_device), want, k, d)


def test_noalloc_rejects_sizes_beyond_its_fixed_scratch(pkg, cuda_device):
    import torch

    lib = pkg._lib.load()
    buf = torch.zeros(64, dtype=torch.float64, device=cuda_device)
    p = ctypes.c_void_p(buf.data_ptr())
    rc = lib.kmc_cluster_sums_noalloc(0, p, p, p, (1 << 20) + 1, 2, 4, p, None)
    assert rc == -3 and b"at most 1048576 rows" in lib.kmc_last_error()
    rc = lib.kmc_cluster_sums_noalloc(0, p, p, p, 1000, 2, 4097, p, None)
    assert rc == -3 and b"at most 4096 clusters" in lib.kmc_last_error()


def _numpy_partials(X, labels, C):
    """Per-cluster float64 sums, counts and inertia of the rows whose label lies in [0, K)."""
    K, D = C.shape
    ok = (labels >= 0) & (labels < K)
    Xo, lo = X[ok].astype(np.float64), labels[ok]
    sums = np.zeros((K, D))
    np.add.at(sums, lo, Xo)
    counts = np.bincount(lo, minlength=K).astype(np.float64)
    inertia = float(((Xo - C.astype(np.float64)[lo]) ** 2).sum())
    return sums, counts, inertia


@pytest.mark.parametrize("n,d,k,bad", [(1 << 20, 8, 4096, 0), (50_001, 7, 33, 0), (70_000, 16, 100, 5)])
def test_noalloc_against_numpy_at_the_limits_and_on_odd_shapes(n, d, k, bad, pkg, cuda_device):
    """Checked straight against NumPy: exactly 1 048 576 rows and 4096 clusters (the limits of the fixed scratch),
    a row length of 7 floats (no 16-byte loads possible), and labels outside [0, K) (those rows are left out)."""
    import torch

    rng = np.random.default_rng(n + d)
    X = rng.normal(0.0, 3.0, size=(n, d)).astype(np.float32)
    C = rng.normal(0.0, 3.0, size=(k, d)).astype(np.float32)
    labels = rng.integers(0, k, size=n).astype(np.int32)
    if bad:
        labels[rng.choice(n, bad, replace=False)] = np.array([-1, k, k + 7, -100, 2**30][:bad], dtype=np.int32)
    pts, lab, cent = (torch.from_numpy(a).to(cuda_device) for a in (X, labels, C))
    for _ in range(2):                                     # twice in a row on the shared scratch
        got = _noalloc_partials(pkg, pts, lab, cent, cuda_device).cpu().numpy()
        sums, counts, inertia = _numpy_partials(X, labels, C)
        assert np.array_equal(got[k * d:k * d + k], counts) and counts.sum() == n - bad
        assert np.max(np.abs(got[:k * d].reshape(k, d) - sums)) <= 1e-9 * max(1.0, np.max(np.abs(sums)))
        assert abs(got[k * d + k] - inertia) <= RTOL * inertia
    assert torch.cuda.current_device() == (cuda_device.index or 0)
