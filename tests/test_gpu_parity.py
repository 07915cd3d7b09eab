"""GPU parity tests (B200): the CUDA path, called through the C ABI (ctypes -> libkmc_b200.so),
against the float64 oracle on the same seeded inputs and against the committed golden fixtures.

Bar (BASELINE.json north_star): labels identical except rows whose oracle top-2 distances differ
by <= 1e-5 relative; centroids and inertia within 1e-4 relative at every iteration; counts are
integers and must match exactly once near-tie rows are set aside.
"""
import numpy as np
import pytest

import oracle
from conftest import load_golden

pytestmark = pytest.mark.gpu

RTOL = 1e-4          # centroids / inertia, per iteration (north_star)
TIE_RTOL = 1e-5      # near-tie window for labels (north_star)


def _shard(pkg, X, K, device, path="auto", empty="keep"):
    import torch

    s = pkg.LloydShard(X.shape[0], X.shape[1], K, device, empty=empty, path=path)
    s.bind_points(torch.from_numpy(np.ascontiguousarray(X, dtype=np.float32)).to(device))
    return s


def _check_iterations(pkg, X, C0, iters, device, path="auto", n_shards=1, form="diff", hist=None, tc_terms=None,
                      params=()):
    """Run `iters` Lloyd iterations on the GPU, one C-ABI call per phase, checking every
    intermediate against the oracle (labels_t, counts_t, inertia_t vs C_t; C_{t+1})."""
    import torch

    K = C0.shape[0]
    ranges = pkg.shard_ranges(X.shape[0], n_shards)
    shards = [_shard(pkg, X[r0:r1], K, device, path) for r0, r1 in ranges]
    for s in shards:
        s.set_centroids(C0)
        if tc_terms is not None:
            s.set_param(pkg._lib.KMC_PARAM_TC_TERMS, tc_terms)
        for pid, val in params:
            s.set_param(pid, val)
    C = np.asarray(C0, dtype=np.float64)
    for t in range(iters):
        ref = oracle.lloyd_iteration(X, C, n_shards=n_shards, form=form) if hist is None else None
        if hist is not None:
            lab_ref, d1, d2 = oracle.assign(X, C, form)
            ref = {"labels": lab_ref, "d1": d1, "d2": d2, "counts": hist["counts"][t],
                   "centroids": hist["centroids"][t], "inertia": hist["inertia"][t]}
            assert np.array_equal(lab_ref, hist["labels"][t])
        labels = np.concatenate([s.assign().cpu().numpy() for s in shards])
        bad = oracle.label_mismatches(labels, ref["labels"], ref["d1"], ref["d2"], TIE_RTOL)
        assert len(bad) == 0, f"iter {t}: {len(bad)} label mismatches outside near-ties, first rows {bad[:5]}"
        for s in shards:
            s.accumulate()
        total = torch.zeros_like(shards[0].partials)
        for s in shards:
            total += s.partials
        for s in shards:
            s.partials.copy_(total)             # what the all-reduce leaves in every shard
        part = total.cpu().numpy()
        counts, inertia = part[K * X.shape[1]:K * X.shape[1] + K], part[K * X.shape[1] + K]
        n_tie = int(oracle.near_tie_mask(ref["d1"], ref["d2"], TIE_RTOL).sum())
        assert counts.sum() == X.shape[0]
        assert np.abs(counts - ref["counts"]).sum() <= 2 * n_tie, f"iter {t}: counts differ beyond near-ties"
        assert abs(inertia - ref["inertia"]) <= RTOL * abs(ref["inertia"]), (t, inertia, ref["inertia"])
        for s in shards:
            s.update()
        got = shards[0].get_centroids()
        for s in shards[1:]:
            assert np.array_equal(s.get_centroids(), got), "replicated centroids diverged"
        assert oracle.rel_err(got, ref["centroids"]) <= RTOL, (t, oracle.rel_err(got, ref["centroids"]))
        C = np.asarray(ref["centroids"], dtype=np.float64)
    stats = shards[0].stats()
    for s in shards:
        s.close()
    return stats


GOLDEN_CASES = ["c1_blobs_d2_k8", "c1_blobs_d2_k8_g3", "c2_image_d3_k16", "c3_blobs_d32_k256",
                "c4_blobs_d128_k1024", "c5_blobs_d64_k4096", "edge_k1_d1", "edge_dup_centroids"]


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_golden_case_per_iteration(name, pkg, golden_index, cuda_device):
    z, meta = load_golden(name, pkg, golden_index)
    stats = _check_iterations(pkg, z["X"], z["C0"], meta["iters"], cuda_device, n_shards=meta["shards"],
                              form=meta["form"], hist=z)
    want = {"c1": "small", "c2": "small", "c3": "tc", "c4": "tc", "c5": "tc"}.get(name[:2])
    if want:
        assert stats["path"] == want


@pytest.mark.parametrize("n,d,k,path", [
    (5000, 2, 8, "small"), (4097, 3, 16, "small"), (3001, 8, 40, "small"), (1, 2, 1, "small"),
    (2003, 2, 100, "small"), (1501, 1, 150, "small"), (2500, 5, 64, "small"), (70001, 3, 16, "small"),
    (3000, 5, 7, "simt"), (2500, 20, 100, "simt"), (2049, 33, 65, "simt"), (4096, 32, 256, "simt"),
    (4000, 16, 32, "tc"), (3000, 32, 256, "tc"), (2500, 64, 300, "tc"), (2000, 128, 1024, "tc"),
    (1500, 100, 257, "tc"), (1200, 160, 96, "tc"), (1000, 256, 512, "tc"), (129, 64, 64, "tc"),
])
def test_random_shapes_each_path(n, d, k, path, pkg, cuda_device):
    X, _ = pkg.datagen.blobs_numpy(n, d, max(2, k // 3), seed=n + d + k)
    stats = _check_iterations(pkg, X, X[:k].copy(), 2, cuda_device, path=path,
                              form="gemm" if d * k > 20000 else "diff")
    assert stats["path"] == path


@pytest.mark.parametrize("terms", [1, 3, 16])
@pytest.mark.parametrize("n,d,k", [(6000, 128, 1024), (5000, 64, 300), (3000, 256, 512), (2500, 32, 256)])
def test_tensor_path_with_one_and_three_mma_terms(n, d, k, terms, pkg, cuda_device):
    """Every precision tier of the tensor kernel must end in the same exact labels: with one MMA term (fp32 or,
    terms = 16, fp16 accumulators) many more rows go through the exact refine kernel, with three almost none."""
    X, _ = pkg.datagen.blobs_numpy(n, d, max(2, k // 4), seed=n + d + k)      # several centroids per blob
    stats = _check_iterations(pkg, X, X[:k].copy(), 2, cuda_device, path="tc", form="gemm", tc_terms=terms)
    assert stats["path"] == "tc"


@pytest.mark.parametrize("chunks,terms", [(4, 1), (3, 3), (2, 0), (3, 16)])
def test_chunk_pipelined_accumulate(chunks, terms, pkg, cuda_device):
    """KMC_PARAM_CHUNKS > 1: the shard is assigned chunk by chunk while the finished chunk is refined on a
    second stream; check every iteration against the oracle, with and without the incremental accumulation."""
    n, d, k = 50_000 + 77, 128, 512
    X, _ = pkg.datagen.blobs_numpy(n, d, k // 4, seed=99)
    for inc in (1, 0):
        stats = _check_iterations(pkg, X, X[:k].copy(), 3, cuda_device, path="tc", form="gemm", tc_terms=terms,
                                  params=[(pkg._lib.KMC_PARAM_CHUNKS, chunks), (pkg._lib.KMC_PARAM_INCREMENTAL, inc)])
        assert stats["path"] == "tc"


def test_tensor_fallback_on_poorly_separated_data(pkg, cuda_device):
    """Points far from the origin with tiny spread: one fp16 MMA term cannot certify anything, so every row goes
    through the exact refine (bounded walks, the tile kernel behind them) or, once the walks are known to be
    long, through the device-side fallback to the 3-term kernel -- same exact labels either way."""
    # (one iteration per start: with coordinates near 1000 and a spread of 0.05 the fp32 rounding of the
    # UPDATED centroids already moves distances by more than the 1e-5 near-tie window, so later iterations
    # are not comparable with the float64 oracle on this data set)
    for seed in (3, 4):
        rng = np.random.default_rng(seed)
        n, d, k = 20000, 64, 128
        X = (1000.0 + rng.normal(0.0, 0.05, size=(n, d))).astype(np.float32)
        _check_iterations(pkg, X, X[:k].copy(), 1, cuda_device, path="tc", form="diff")


def test_unscaled_data_ranges_on_tensor_path(pkg, cuda_device):
    # tiny and huge magnitudes go through the power-of-two scaling of the fp16 planes
    for scale in (1e-6, 3e4):
        X, _ = pkg.datagen.blobs_numpy(3000, 64, 40, seed=17)
        X = (X * np.float32(scale)).astype(np.float32)
        _check_iterations(pkg, X, X[:128].copy(), 2, cuda_device, path="tc")


def test_tensor_scores_match_fp64(pkg, cuda_device):
    """The raw tcgen05 scores |c|^2 - 2 x.c against float64, within the bound the kernel uses to
    decide which rows need the exact refine pass."""
    for n, d, k in [(300, 64, 256), (1000, 128, 1024), (257, 32, 100), (640, 256, 300)]:
        X, _ = pkg.datagen.blobs_numpy(n, d, 16, seed=d)
        C = X[:k].copy() if k <= n else np.concatenate([X] * (k // n + 1))[:k] * np.float32(0.9)
        s = _shard(pkg, X, k, cuda_device, path="tc")
        s.set_centroids(C)
        got = s.debug_tc_scores().cpu().numpy().astype(np.float64)
        X64, C64 = X.astype(np.float64), C.astype(np.float64)
        want = (C64 * C64).sum(1)[None, :] - 2.0 * X64 @ C64.T
        bound = np.linalg.norm(X64, axis=1)[:, None] * (2 * np.linalg.norm(C64, axis=1))[None, :]
        err = np.abs(got - want) / bound
        d_pad = {32: 64, 64: 64, 128: 128, 256: 256}[d]
        gamma = 1.25 * (3 * d_pad / 16 + 6) / 2 ** 23       # the bound kmc_create derives (kmc_api.cu)
        print(f"tc scores n={n} d={d} k={k}: max rel err {err.max():.3e} (gamma {gamma:.3e})")
        assert err.max() < gamma, (n, d, k, err.max())
        s.close()


def test_exact_ties_resolve_to_lowest_index_on_every_path(pkg, cuda_device):
    for d, k, path in [(4, 6, "small"), (20, 12, "simt"), (32, 64, "tc"), (128, 300, "tc")]:
        X, _ = pkg.datagen.blobs_numpy(2000, d, 5, seed=d)
        C = X[:k].copy()
        C[k - 1] = C[0]                      # duplicate centroid: rows nearest to it tie exactly
        C[3] = C[2]
        s = _shard(pkg, X, k, cuda_device, path=path)
        s.set_centroids(C)
        lab = s.assign().cpu().numpy()
        ref, d1, d2 = oracle.assign(X, C)
        assert (ref != k - 1).all() and (ref != 3).all() and (d2 == d1).sum() > 0
        assert np.array_equal(lab[d2 == d1], ref[d2 == d1])              # exact ties: lowest index
        assert len(oracle.label_mismatches(lab, ref, d1, d2, TIE_RTOL)) == 0
        s.close()


def test_empty_cluster_policies(pkg, cuda_device):
    X = np.zeros((64, 2), dtype=np.float32)
    X[:, 0] = np.arange(64) % 2
    C0 = np.array([[0.0, 0.0], [1.0, 0.0], [50.0, 50.0]], dtype=np.float32)
    for empty in ("keep", "nan"):
        s = _shard(pkg, X, 3, cuda_device, empty=empty)
        s.set_centroids(C0)
        s.iterate(2)
        C = s.get_centroids()
        assert np.allclose(C[:2], [[0, 0], [1, 0]])
        assert (np.isnan(C[2]).all() if empty == "nan" else np.array_equal(C[2], C0[2]))
        s.close()


def test_nan_centroid_never_wins(pkg, cuda_device):
    for d, k, path in [(3, 8, "small"), (24, 16, "simt"), (64, 64, "tc")]:
        X, _ = pkg.datagen.blobs_numpy(1500, d, 4, seed=2)
        C = X[:k].copy()
        C[1] = np.nan
        s = _shard(pkg, X, k, cuda_device, path=path)
        s.set_centroids(C)
        lab = s.assign().cpu().numpy()
        ref, d1, d2 = oracle.assign(X, C)
        assert (lab != 1).all()
        assert len(oracle.label_mismatches(lab, ref, d1, d2, TIE_RTOL)) == 0
        s.close()


def test_host_staging_fp64_fp32_and_batches(pkg, cuda_device):
    X, _ = pkg.datagen.blobs_numpy(10007, 7, 5, seed=1)
    for host in (X, X.astype(np.float64)):
        for batch in (None, 1000, 4096):
            s = pkg.LloydShard(5000, 7, 5, cuda_device)
            s.stage_from_host(host, 3000, batch)
            assert np.array_equal(s.points.cpu().numpy(), X[3000:8000])
            s.close()


@pytest.mark.parametrize("terms", [1, 3, 16])
def test_well_separated_rows_are_certified_by_the_tensor_kernel(terms, pkg, cuda_device):
    """Guards the epilogue's best / second-best bookkeeping: on blobs with one centroid each nearly every row
    must be certified by the tensor kernel itself (a wrong second-best would flood or starve the refine)."""
    k = 256
    X, centers = pkg.datagen.blobs_numpy(40000, 128, k, seed=21)
    s = _shard(pkg, X, k, cuda_device, path="tc")
    s.set_centroids(centers.astype(np.float32))
    s.set_param(pkg._lib.KMC_PARAM_TC_TERMS, terms)
    lab = s.assign().cpu().numpy()
    ref, d1, d2 = oracle.assign(X, centers.astype(np.float32))
    assert len(oracle.label_mismatches(lab, ref, d1, d2, TIE_RTOL)) == 0
    assert s.stats()["ambiguous"] <= 0.002 * X.shape[0], s.stats()
    s.close()


@pytest.mark.parametrize("kernel", [0, 4, 8])
def test_refine_kernels_agree_with_the_oracle_on_long_and_short_walks(kernel, pkg, cuda_device):
    """Both exact-refine kernels (warp per row / 8 lanes per row) and the on-device choice between them:
    iteration 1 from C0 = X[0:K] has long candidate walks (blobs without a centroid), later ones short."""
    X, _ = pkg.datagen.blobs_numpy(60000, 64, 96, seed=11)
    C0 = X[:96].copy()
    s = _shard(pkg, X, 96, cuda_device, path="tc")
    s.set_centroids(C0)
    s.set_param(pkg._lib.KMC_PARAM_TC_TERMS, 1)            # every uncertified row goes through the refine
    s.set_param(pkg._lib.KMC_PARAM_REFINE_KERNEL, kernel)
    C = C0.astype(np.float64)
    for it in range(4):
        lab = s.assign().cpu().numpy()
        ref, d1, d2 = oracle.assign(X, C)
        assert len(oracle.label_mismatches(lab, ref, d1, d2, TIE_RTOL)) == 0, f"iteration {it}"
        s.accumulate()
        s.update()
        C = s.get_centroids().astype(np.float64)
    assert s.stats()["ambiguous"] > 0
    s.close()


def test_recolour_is_centre_of_label(pkg, cuda_device):
    X = pkg.datagen.image_numpy(120, 160, seed=7)                      # the colour-quantisation use case
    s = _shard(pkg, X, 16, cuda_device)
    s.set_centroids(X[:16])
    s.iterate(3)
    lab = s.assign()
    got = s.recolour(lab).cpu().numpy()
    assert np.array_equal(got, s.get_centroids()[lab.cpu().numpy()])
    assert np.array_equal(s.recolour().cpu().numpy(), got)
    s.close()


def test_labels_to_host_int32_int64_pageable_and_pinned(pkg, cuda_device):
    """kmc_labels_to_host: several staging chunks (> 16 MB of labels), both widths, pageable and pinned targets."""
    import torch

    n = 5_000_011                                    # 40 MB of int64 labels: three chunks, ragged tail
    X, _ = pkg.datagen.blobs_numpy(n, 2, 8, seed=5)
    s = _shard(pkg, X, 8, cuda_device)
    s.set_centroids(X[:8])
    dev = s.assign()
    want = dev.cpu().numpy()
    for dt in (np.int32, np.int64):
        out = s.labels_to_host(dev, np.full(n, -7, dtype=dt))
        assert out.dtype == dt and np.array_equal(out, want)
        pinned = torch.empty(n, dtype=torch.int32 if dt == np.int32 else torch.int64, pin_memory=True)
        s.labels_to_host(dev, pinned.numpy())
        assert np.array_equal(pinned.numpy(), want)
    with pytest.raises(ValueError):
        s.labels_to_host(dev, np.empty(n, dtype=np.int16))
    s.close()


def test_streamed_first_iteration_equals_the_plain_path(pkg, cuda_device):
    """kmc_bind_points_streamed + kmc_accumulate_rows: rows arriving in pieces give the sums of one plain
    accumulate over all rows, and the shard continues like a normally bound one."""
    n, d, k = 150_000, 64, 64
    X, _ = pkg.datagen.blobs_numpy(n, d, 40, seed=9)
    C0 = X[:k].copy()
    plain = _shard(pkg, X, k, cuda_device, path="tc")
    plain.set_centroids(C0)
    plain.accumulate()
    want = plain.partials.cpu().numpy().copy()
    for host in (X, X.astype(np.float64)):
        s = pkg.LloydShard(n, d, k, cuda_device, path="tc")
        s.set_centroids(C0)
        with pytest.raises(pkg._lib.KmcError):
            s.accumulate()                                       # nothing bound yet
        assert s.stage_and_accumulate_from_host(host, 0, piece_rows=20_000)      # 8 pieces, ragged tail
        got = s.partials.cpu().numpy()
        assert np.array_equal(got[k * d:k * d + k], want[k * d:k * d + k])        # counts
        assert oracle.rel_err(got[:k * d], want[:k * d]) <= 1e-6      # fp32 lane sums over different row runs
        assert abs(got[k * d + k] - want[k * d + k]) <= 1e-6 * want[k * d + k]     # inertia
        s.update()
        s.iterate(2)
        plain2 = _shard(pkg, X, k, cuda_device, path="tc")
        plain2.set_centroids(C0)
        plain2.iterate(3)
        assert oracle.rel_err(s.get_centroids(), plain2.get_centroids()) <= 1e-6
        plain2.close()
        s.close()
    plain.close()
    # small-D shards do not stream
    s = pkg.LloydShard(100_000, 2, 8, cuda_device)
    s.set_centroids(X[:8, :2].copy())
    assert s.stage_and_accumulate_from_host(np.ascontiguousarray(X[:100_000, :2]), 0) is False
    s.close()


def test_streamed_load_falls_back_when_later_rows_outgrow_the_plane_scale(pkg, cuda_device):
    n, d, k = 120_000, 64, 64
    X, _ = pkg.datagen.blobs_numpy(n, d, 40, seed=10)
    X[:30_000] *= 1e-3                       # the first piece fixes a scale 1000x too large for the rest
    C0 = X[40_000:40_000 + k].copy()
    s = pkg.LloydShard(n, d, k, cuda_device, path="tc")
    s.set_centroids(C0)
    assert s.stage_and_accumulate_from_host(X, 0, piece_rows=30_000)
    assert s.stats()["x_scaled_max"] < 32768.0                   # rebound at the global scale
    plain = _shard(pkg, X, k, cuda_device, path="tc")
    plain.set_centroids(C0)
    plain.accumulate()
    got, want = s.partials.cpu().numpy(), plain.partials.cpu().numpy()
    assert np.array_equal(got[k * d:k * d + k], want[k * d:k * d + k]) and oracle.rel_err(got[:k * d], want[:k * d]) <= 1e-6
    s.close()
    plain.close()


def test_drop_in_streams_the_first_iteration_on_large_tensor_shards(pkg, cuda_device):
    n, d, k = 100_000, 64, 64
    X, _ = pkg.datagen.blobs_numpy(n, d, 40, seed=12)
    C0 = X[:k].copy()
    ref = oracle.run_lloyd(X, C0, 3)
    for stream in (True, False):
        res = pkg.distribuited_k_means(X, k, 1, C0, 3, stream_first_iteration=stream)
        assert res["n_iter"] == 3
        assert oracle.rel_err(res["end_center"], ref["end_center"]) <= RTOL
        assert len(oracle.label_mismatches(res["cluster_idx"], ref["cluster_idx"], ref["final_d1"], ref["final_d2"])) == 0
        assert abs(res["inertia"] - ref["history"][-1]["inertia"]) <= RTOL * ref["history"][-1]["inertia"]
    one = pkg.distribuited_k_means(X, k, 1, C0, 1)
    assert one["n_iter"] == 1 and abs(one["inertia"] - ref["history"][0]["inertia"]) <= RTOL * ref["history"][0]["inertia"]


def test_drop_in_entry_point_matches_oracle(pkg, cuda_device):
    X, _ = pkg.datagen.blobs_numpy(20000, 2, 8, seed=123128)
    X64 = X.astype(np.float64)                       # the reference feeds float64 (make_classification)
    C0 = X64[:8]
    res = pkg.distribuited_k_means(X64, 8, ["/device:GPU:0"], C0, 20)          # 5-positional form (:303)
    ref = oracle.run_lloyd(X, C0, 20)
    for key in ("end_center", "cluster_idx", "centroids", "init_center", "setup_time", "initialization_time",
                "computation_time", "n_iter"):
        assert key in res
    assert res["n_iter"] == 20 and res["cluster_idx"].dtype == np.int64
    assert oracle.rel_err(res["end_center"], ref["end_center"]) <= RTOL
    assert len(oracle.label_mismatches(res["cluster_idx"], ref["cluster_idx"], ref["final_d1"], ref["final_d2"])) == 0
    assert abs(res["inertia"] - ref["history"][-1]["inertia"]) <= RTOL * ref["history"][-1]["inertia"]
    # `centroids` (:279, :286): labels of the final iteration against the centroids BEFORE its update
    assert res["centroids"].dtype == np.int64 and res["centroids"].shape == (20000,)
    last = ref["history"][-1]
    assert len(oracle.label_mismatches(res["centroids"], ref["centroids"], last["d1"], last["d2"])) == 0
    # ... which differs from cluster_idx while labels still move: two iterations from a poor start
    res2 = pkg.distribuited_k_means(X64, 8, ["/device:GPU:0"], C0, 2)
    ref2 = oracle.run_lloyd(X, C0, 2)
    assert (ref2["centroids"] != ref2["cluster_idx"]).sum() > 0
    last2 = ref2["history"][-1]
    assert len(oracle.label_mismatches(res2["centroids"], ref2["centroids"], last2["d1"], last2["d2"])) == 0
    assert len(oracle.label_mismatches(res2["cluster_idx"], ref2["cluster_idx"], ref2["final_d1"], ref2["final_d2"])) == 0
    # HEAD's 4-positional form (:180): centres from seeded k-means++
    res4 = pkg.distribuited_k_means(X, 8, 1, 5, seed=3)
    assert res4["n_iter"] == 5 and res4["init_center"].shape == (8, 2)
    ref4 = oracle.run_lloyd(X, res4["init_center"], 5)
    assert oracle.rel_err(res4["end_center"], ref4["end_center"]) <= RTOL
    # tolerance-based early stop (this project's addition)
    res_tol = pkg.distribuited_k_means(X, 8, 1, C0, 200, tol=1e-12)
    assert res_tol["converged"] and res_tol["n_iter"] < 200


def test_size_independent_properties_large(pkg, cuda_device):
    """A larger tensor-path run checked through properties that need no full oracle pass:
    counts sum to N, the per-cluster sums add up to the column sums of X (linearity), the label
    pass is idempotent, and a random row sample agrees with the oracle."""
    import torch

    n, d, k = 1 << 20, 128, 1024
    X, _ = pkg.datagen.blobs_torch(n, d, k, seed=123128, device=cuda_device)
    C0 = X[:k].cpu().numpy()
    s = pkg.LloydShard(n, d, k, cuda_device)
    s.bind_points(X)
    s.set_centroids(C0)
    lab1 = s.assign().clone()
    lab2 = s.assign()
    assert torch.equal(lab1, lab2)
    s.accumulate()
    part = s.partials.cpu().numpy()
    sums, counts = part[:k * d].reshape(k, d), part[k * d:k * d + k]
    assert counts.sum() == n
    assert np.array_equal(counts, np.bincount(lab1.cpu().numpy(), minlength=k))
    col = X.double().sum(0).cpu().numpy()
    assert np.max(np.abs(sums.sum(0) - col)) <= 1e-6 * np.max(np.abs(col))
    idx = np.random.default_rng(0).choice(n, 4096, replace=False)
    ref, d1, d2 = oracle.assign(X[torch.from_numpy(idx).to(cuda_device)].cpu().numpy(), C0, "gemm")
    assert len(oracle.label_mismatches(lab1.cpu().numpy()[idx], ref, d1, d2, TIE_RTOL)) == 0
    assert abs(part[k * d + k] - d1.sum() * (n / 4096)) < 0.05 * part[k * d + k]
    s.close()


@pytest.mark.parametrize("n,d,k,batch", [(10007, 2, 8, 3000), (9001, 64, 128, 4096)])
def test_out_of_core_batches_are_one_exact_lloyd_run(n, d, k, batch, pkg, cuda_device):
    """Rows kept in host memory and streamed in batches every iteration (kmc_resize /
    kmc_accumulate_add) must give the same run as resident rows: batches add up to ONE Lloyd iteration."""
    X, _ = pkg.datagen.blobs_numpy(n, d, k, seed=77)
    C0 = X[:k].copy()
    res = pkg.distribuited_k_means(X, k, 1, C0, 5, batch_size=batch, out_of_core=True)
    assert res["out_of_core"] and res["batch_rows"] == [batch]
    ref = oracle.run_lloyd(X, C0, 5, form="gemm" if d * k > 4096 else "diff")
    assert oracle.rel_err(res["end_center"], ref["end_center"]) <= RTOL
    assert len(oracle.label_mismatches(res["cluster_idx"], ref["cluster_idx"], ref["final_d1"], ref["final_d2"])) == 0
    assert abs(res["inertia"] - ref["history"][-1]["inertia"]) <= RTOL * ref["history"][-1]["inertia"]
    resident = pkg.distribuited_k_means(X, k, 1, C0, 5)
    assert oracle.rel_err(res["end_center"], resident["end_center"]) <= 1e-6


def test_device_kmeanspp_matches_its_cpu_restatement(pkg, cuda_device):
    """kmc_kmeanspp against oracle/kmeanspp_oracle.py on integer-valued data (every squared distance is
    exact in float32, so the device must choose exactly the same rows), plus the seeded 4-positional
    form of the entry point (reference :180, :191)."""
    import torch
    from oracle.kmeanspp_oracle import kmeanspp

    rng = np.random.default_rng(5)
    for n, d, k in [(5000, 3, 16), (2500, 17, 40), (1100, 2, 8)]:
        X = rng.integers(-20, 21, size=(n, d)).astype(np.float32)
        for seed in (1, 123128):
            want_c, want_rows = kmeanspp(X, k, seed)
            got_c, got_rows = pkg.kmeanspp_device(torch.from_numpy(X).to(cuda_device), k, seed)
            assert np.array_equal(got_rows, want_rows), (n, d, k, seed)
            assert np.array_equal(got_c, want_c)
    # through the entry point: same seed -> same centres, centres are rows of the data
    X, _ = pkg.datagen.blobs_numpy(20000, 2, 8, seed=2)
    a = pkg.distribuited_k_means(X, 8, 1, 3, seed=7)
    b = pkg.distribuited_k_means(X, 8, 1, 3, seed=7)
    assert np.array_equal(a["init_center"], b["init_center"]) and np.array_equal(a["end_center"], b["end_center"])
    assert all((X == c).all(1).any() for c in a["init_center"])
    ref = oracle.run_lloyd(X, a["init_center"], 3)
    assert oracle.rel_err(a["end_center"], ref["end_center"]) <= RTOL


@pytest.mark.parametrize("n,d,k,path", [(60_000, 64, 96, "tc"), (40_000, 128, 256, "tc"), (30_000, 20, 50, "simt"),
                                         (9_000, 33, 65, "simt")])
def test_incremental_accumulation_equals_full_passes(n, d, k, path, pkg, cuda_device):
    """kmc_accumulate with running sums (changed rows only) against KMC_PARAM_INCREMENTAL = 0 (every pass sums
    every row), iteration by iteration from C0 = X[0:K]: identical counts, sums to 1e-12, inertia to 1e-7, and
    the number of changed rows the library reports equals the number of labels that moved."""
    X, _ = pkg.datagen.blobs_numpy(n, d, max(2, k // 3), seed=n + d)
    C0 = X[:k].copy()
    inc = _shard(pkg, X, k, cuda_device, path=path)
    full = _shard(pkg, X, k, cuda_device, path=path)
    full.set_param(pkg._lib.KMC_PARAM_INCREMENTAL, 0)
    for s in (inc, full):
        s.set_centroids(C0)
    prev = None
    saw_incremental = False
    for it in range(8):
        for s in (inc, full):
            s.accumulate()
        a, b = inc.partials.cpu().numpy(), full.partials.cpu().numpy()
        assert np.array_equal(a[k * d:k * d + k], b[k * d:k * d + k]), f"iteration {it}: counts"
        assert oracle.rel_err(a[:k * d], b[:k * d]) <= 1e-12, (it, oracle.rel_err(a[:k * d], b[:k * d]))
        assert abs(a[k * d + k] - b[k * d + k]) <= 1e-7 * b[k * d + k], (it, a[k * d + k], b[k * d + k])
        lab = inc.last_labels().cpu().numpy()
        assert np.array_equal(lab, full.last_labels().cpu().numpy())
        st = inc.stats()
        if prev is not None:
            moved = int((lab != prev).sum())
            assert st["changed_rows"] == moved, (it, st, moved)
            assert st["full_pass"] == (moved > n // 4)
            saw_incremental |= not st["full_pass"]
        else:
            assert st["full_pass"] and full.stats()["full_pass"]
        prev = lab
        for s in (inc, full):
            s.update()
        assert np.array_equal(inc.get_centroids(), full.get_centroids()) or \
            oracle.rel_err(inc.get_centroids(), full.get_centroids()) <= 1e-7
    assert saw_incremental
    inc.close()
    full.close()


def test_incremental_accumulation_falls_back_when_most_labels_move(pkg, cuda_device):
    """More than a quarter of the rows changing their label (new centroids set from outside) must make the
    device take the full pass, and a later quiet iteration the incremental one again."""
    n, d, k = 50_000, 64, 64
    X, _ = pkg.datagen.blobs_numpy(n, d, 40, seed=31)
    s = _shard(pkg, X, k, cuda_device, path="tc")
    s.set_centroids(X[:k].copy())
    s.accumulate(); s.update()
    s.accumulate()
    assert not s.stats()["full_pass"]
    s.set_centroids(X[1000:1000 + k][::-1].copy())            # unrelated centroids: nearly every label changes
    s.accumulate()
    st = s.stats()
    assert st["full_pass"] and st["changed_rows"] > n // 4
    part = s.partials.cpu().numpy()
    ref = oracle.lloyd_iteration(X, X[1000:1000 + k][::-1].astype(np.float64), form="diff")
    assert np.abs(part[k * d:k * d + k] - ref["counts"]).sum() <= 2 * int(oracle.near_tie_mask(ref["d1"], ref["d2"], TIE_RTOL).sum())
    assert abs(part[k * d + k] - ref["inertia"]) <= RTOL * ref["inertia"]
    s.update()
    s.accumulate()
    assert not s.stats()["full_pass"]
    s.close()


def test_incremental_inertia_on_data_far_from_the_origin(pkg, cuda_device):
    """The incremental pass derives the inertia from the sums (sum|x|^2 - 2 c.S + n|c|^2): with coordinates near
    1000 and a spread of 0.5 that difference cancels nine digits, which the fp64 sums must carry."""
    rng = np.random.default_rng(8)
    n, d, k = 30_000, 32, 40
    centers = 1000.0 + rng.normal(0.0, 4.0, size=(k, d))
    X = (centers[rng.integers(0, k, n)] + rng.normal(0.0, 0.5, size=(n, d))).astype(np.float32)
    s = _shard(pkg, X, k, cuda_device, path="simt")
    s.set_centroids(X[:k].copy())
    C = X[:k].astype(np.float64)
    for it in range(5):
        s.accumulate()
        ref = oracle.lloyd_iteration(X, C, form="diff")
        got = s.partials.cpu().numpy()[k * d + k]
        assert abs(got - ref["inertia"]) <= RTOL * ref["inertia"], (it, got, ref["inertia"], s.stats())
        s.update()
        C = s.get_centroids().astype(np.float64)
    assert not s.stats()["full_pass"]
    s.close()


def test_last_labels_on_every_path(pkg, cuda_device):
    """kmc_last_labels: the labels of the last accumulate pass (against the pre-update centroids) on the small-D
    path (recomputed from the saved centroids), the SIMT path and the tensor path."""
    for n, d, k, path in [(20_000, 3, 16, "small"), (8_000, 20, 30, "simt"), (20_000, 64, 64, "tc")]:
        X, _ = pkg.datagen.blobs_numpy(n, d, max(2, k // 2), seed=d)
        C0 = X[:k].copy()
        s = _shard(pkg, X, k, cuda_device, path=path)
        s.set_centroids(C0)
        C = C0.astype(np.float64)
        for it in range(2):
            s.accumulate()
            ref, d1, d2 = oracle.assign(X, C)
            before = s.last_labels().cpu().numpy()                  # before the update
            assert len(oracle.label_mismatches(before, ref, d1, d2, TIE_RTOL)) == 0, (path, it)
            s.update()
            after = s.last_labels().cpu().numpy()                   # after it: still the same pass
            assert np.array_equal(before, after), (path, it)
            C = s.get_centroids().astype(np.float64)
        s.close()


def test_gpu_mirror_matches_the_numpy_oracle(pkg, cuda_device):
    """oracle/gpu_mirror.py (float64 torch on the GPU, used for the full-size checks of bench.py) against the
    NumPy oracle it restates: identical labels outside near-ties, distances, sums, counts and inertia."""
    import torch
    from oracle import gpu_mirror

    for n, d, k, form in [(20_000, 2, 8, "diff"), (9_000, 32, 256, "diff"), (5_000, 128, 1024, "gemm"), (300, 3, 1, "diff")]:
        X, _ = pkg.datagen.blobs_numpy(n, d, max(1, k // 2), seed=n)
        C = X[:k].astype(np.float64)
        if k > 2:
            C[2] = C[1]                                              # an exact tie: lowest index must win
        ref = oracle.lloyd_iteration(X, C, form=form)
        got = gpu_mirror.lloyd_pass_f64(torch.from_numpy(X).to(cuda_device), torch.from_numpy(C).to(cuda_device), chunk_rows=4096)
        lab = got["labels"].cpu().numpy()
        assert len(oracle.label_mismatches(lab, ref["labels"], ref["d1"], ref["d2"], TIE_RTOL)) == 0
        exact_tie = ref["d1"] == ref["d2"]
        assert np.array_equal(lab[exact_tie], ref["labels"][exact_tie])
        assert np.allclose(got["d1"].cpu().numpy(), ref["d1"], rtol=1e-9, atol=1e-9 * float(ref["d1"].max()))
        same = lab == ref["labels"]
        assert oracle.rel_err(got["sums"].cpu().numpy(), ref["sums"]) <= 1e-6 + 1e-3 * (~same).mean()
        assert np.abs(got["counts"].cpu().numpy() - ref["counts"]).sum() <= 2 * (~same).sum()
        assert abs(float(got["inertia"]) - ref["inertia"]) <= 1e-9 * ref["inertia"]
        cmp = gpu_mirror.compare_pass(got, torch.from_numpy(ref["labels"]).to(cuda_device),
                                      torch.from_numpy(ref["sums"]).to(cuda_device),
                                      torch.from_numpy(ref["counts"]).to(cuda_device), ref["inertia"])
        assert cmp["label_mismatches_outside_near_ties"] == 0 and cmp["inertia_rel_err"] <= 1e-9


@pytest.mark.parametrize("n,d,k,iters", [(50_000, 2, 8, 5), (33_333, 3, 16, 4), (10_001, 8, 40, 1), (7, 1, 3, 2)])
def test_fused_small_path_iterate_equals_accumulate_update(n, d, k, iters, pkg, cuda_device):
    """kmc_iterate on the small-D path runs one fused launch per iteration (the last CTA updates the centroids and
    clears the other partial buffer); it must give what the separate kmc_accumulate / kmc_update calls give, for odd
    and even iteration counts, and leave inertia, shift and last labels of the final iteration behind."""
    X, _ = pkg.datagen.blobs_numpy(n, d, max(2, k // 2), seed=n)
    C0 = X[:k].copy()
    a = _shard(pkg, X, k, cuda_device, path="small")
    b = _shard(pkg, X, k, cuda_device, path="small")
    for s in (a, b):
        s.set_centroids(C0)
    n_iter, inertia, shift2 = a.iterate(iters)
    for _ in range(iters):
        b.accumulate()
        b.update()
    sb = b.stats()
    assert n_iter == iters
    assert oracle.rel_err(a.get_centroids(), b.get_centroids()) <= 1e-6
    assert abs(inertia - sb["inertia"]) <= 1e-6 * sb["inertia"] and abs(shift2 - sb["shift2"]) <= 1e-6 * max(sb["shift2"], 1e-30)
    pa, pb = a.partials.cpu().numpy(), b.partials.cpu().numpy()
    assert np.array_equal(pa[k * d:k * d + k], pb[k * d:k * d + k]) and pa[k * d:k * d + k].sum() == n
    assert np.array_equal(a.last_labels().cpu().numpy(), b.last_labels().cpu().numpy())
    ref = oracle.run_lloyd(X, C0, iters)
    ties = sum(int(oracle.near_tie_mask(h["d1"], h["d2"], TIE_RTOL).sum()) for h in ref["history"])
    h = ref["history"][-1]
    if iters == 1:
        assert len(oracle.label_mismatches(a.last_labels().cpu().numpy(), h["labels"], h["d1"], h["d2"], TIE_RTOL)) == 0
    if ties == 0:          # (a near-tie row may legitimately go either way and moves a small cluster's centre)
        assert oracle.rel_err(a.get_centroids(), ref["end_center"]) <= RTOL
    assert abs(inertia - ref["history"][-1]["inertia"]) <= RTOL * ref["history"][-1]["inertia"]
    # a second call continues from the state the first one left
    a.iterate(1)
    b.accumulate(); b.update()
    assert oracle.rel_err(a.get_centroids(), b.get_centroids()) <= 1e-6
    a.close()
    b.close()


@pytest.mark.parametrize("n,d,k", [(4096, 128, 1024), (6000, 64, 2048), (3000, 96, 300)])
def test_fp16_accumulator_scores_stay_inside_the_bound_the_kernel_certifies_with(n, d, k, pkg, cuda_device):
    """The fp16-accumulator tier (KMC_PARAM_TC_TERMS = 16) returns S16 |x - c|^2 through packed TMEM reads: the raw
    scores against float64, measured against the accumulation part of the kernel's error bound (one fp16 rounding
    per MMA instruction of the dot product plus the rounding of the final value)."""
    X, _ = pkg.datagen.blobs_numpy(n, d, max(2, k // 2), seed=7)
    C0 = X[:k].copy()
    s = _shard(pkg, X, k, cuda_device, path="tc")
    s.set_centroids(C0)
    s.set_param(pkg._lib.KMC_PARAM_TC_TERMS, 16)
    S = s.debug_tc_scores().double().cpu().numpy()
    s.close()
    X64, C64 = X.astype(np.float64), C0.astype(np.float64)
    T = (X64 * X64).sum(1)[:, None] + (C64 * C64).sum(1)[None, :] - 2.0 * X64 @ C64.T
    xn, cn = np.sqrt((X64 * X64).sum(1)), np.sqrt((C64 * C64).sum(1))
    steps = (-(-d // 64) * 64) // 16
    bound = 1.25 * steps / 2048.0 * 2.0 * xn[:, None] * cn.max() + 1.25 / 2048.0 * np.abs(T)
    assert (np.abs(S - T) / bound).max() < 0.8


def test_fp16_accumulator_tier_with_many_equal_scores(pkg, cuda_device):
    """Integer-valued rows and centroids make many fp16 scores of a row EQUAL (exact ties of the best value in two,
    three and more cells of the epilogue's partition): such rows must reach the refine with a usable anchor and
    end with the exact label (lowest index among equal distances)."""
    rng = np.random.default_rng(5)
    n, d, k = 20000, 64, 1024
    X = rng.integers(-3, 4, size=(n, d)).astype(np.float32)
    C0 = np.unique(X[:4 * k], axis=0)[:k].copy()
    assert C0.shape[0] == k
    C0[1::7] = C0[0::7][: len(C0[1::7])]                       # duplicated centroids: exact ties everywhere
    s = _shard(pkg, X, k, cuda_device, path="tc")
    s.set_centroids(C0)
    s.set_param(pkg._lib.KMC_PARAM_TC_TERMS, 16)
    lab = s.assign().cpu().numpy()
    st = s.stats()
    s.close()
    ref, d1, d2 = oracle.assign(X, C0)
    assert np.array_equal(lab, ref)                             # integer data: distances are exact, ties -> lowest index
    assert st["ambiguous"] > 0
