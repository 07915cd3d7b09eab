"""CPU tests of the oracle: reference sharding rule, the two distance forms, sklearn as an
independent implementation, the reference's NaN artefact and the golden fixtures."""
import numpy as np
import pytest

import oracle
from conftest import load_golden


@pytest.mark.parametrize("n,g", [(10, 4), (10, 1), (7, 7), (5, 8), (0, 3), (1000003, 8)])
def test_shard_sizes_match_array_split(n, g, pkg):
    want = [len(a) for a in np.array_split(np.empty((n, 1)), g)]     # reference :184
    assert oracle.shard_sizes(n, g) == want
    assert pkg.shard_sizes(n, g) == want
    assert sum(want) == n


def test_diff_and_gemm_forms_agree():
    rng = np.random.default_rng(0)
    X = rng.normal(size=(300, 17)).astype(np.float32) * 5 + 20
    C = rng.normal(size=(40, 17)).astype(np.float32) * 5 + 20
    a, b = oracle.dist2(X, C, "diff"), oracle.dist2(X, C, "gemm")
    assert np.max(np.abs(a - b)) / np.max(a) < 1e-12
    la, d1a, d2a = oracle.assign(X, C, "diff")
    lb, d1b, d2b = oracle.assign(X, C, "gemm")
    assert (la == lb).all()


def test_argmin_ties_lowest_index_and_nan_never_wins():
    X = np.array([[0.0, 0.0], [1.0, 1.0]], dtype=np.float32)
    C = np.array([[np.nan, np.nan], [1.0, 0.0], [0.0, 1.0], [1.0, 0.0]], dtype=np.float32)
    lab, d1, d2 = oracle.assign(X, C)
    assert lab.tolist() == [1, 1]            # NaN centroid skipped; tie between 1,2,3 -> lowest
    assert d1.tolist() == [1.0, 1.0] and d2.tolist() == [1.0, 1.0]


@pytest.mark.parametrize("n,d,k,g", [(2000, 2, 8, 1), (1500, 3, 16, 3), (1024, 32, 64, 2), (2048, 128, 96, 8)])
def test_oracle_matches_sklearn_lloyd(n, d, k, g, pkg):
    from sklearn.cluster import KMeans

    X, _ = pkg.datagen.blobs_numpy(n, d, max(k // 2, 1), seed=5)
    C0 = X[:k].copy()
    T = 4
    res = oracle.run_lloyd(X, C0, T, n_shards=g)
    if any((h["counts"] == 0).any() for h in res["history"]):
        pytest.skip("a cluster emptied: sklearn relocates it, the reference does not")
    km = KMeans(n_clusters=k, init=C0.astype(np.float64), n_init=1, algorithm="lloyd", tol=0.0,
                max_iter=T).fit(X.astype(np.float64))
    assert oracle.rel_err(km.cluster_centers_, res["end_center"]) < 1e-10
    assert len(oracle.label_mismatches(km.labels_, res["cluster_idx"], res["final_d1"], res["final_d2"])) == 0


def test_shard_count_does_not_change_lloyd_result(pkg):
    X, _ = pkg.datagen.blobs_numpy(3001, 5, 6, seed=9)
    C0 = X[:6].copy()
    base = oracle.run_lloyd(X, C0, 5, n_shards=1)["end_center"]
    for g in (2, 3, 8):
        assert oracle.rel_err(oracle.run_lloyd(X, C0, 5, n_shards=g)["end_center"], base) < 1e-12


def test_literal_equals_lloyd_when_no_cluster_is_empty(pkg):
    X, _ = pkg.datagen.blobs_numpy(4000, 2, 4, seed=3)
    C = X[:4].astype(np.float64)
    lit = oracle.literal_iteration(X, C, n_shards=2)
    ll = oracle.lloyd_iteration(X, C, n_shards=2)
    assert (lit["counts"] > 0).all()
    assert oracle.rel_err(lit["centroids"], ll["centroids"]) < 1e-12
    assert (lit["labels"] == ll["labels"]).all()


def test_literal_nan_artefact_for_shard_empty_cluster():
    # cluster 1 owns rows only in the second shard: the reference's mean*count gives NaN*0 = NaN
    X = np.array([[0.0], [0.1], [0.2], [10.0]], dtype=np.float64)
    C = np.array([[0.0], [10.0]])
    lit = oracle.literal_iteration(X, C, n_shards=2)
    assert np.isnan(lit["centroids"][1]).all()
    ll = oracle.lloyd_iteration(X, C, n_shards=2)
    assert np.allclose(ll["centroids"], [[0.1], [10.0]])
    assert np.allclose(oracle.literal_iteration(X, C, n_shards=1)["centroids"], [[0.1], [10.0]])


def test_empty_cluster_policies():
    X = np.array([[0.0], [1.0]], dtype=np.float32)
    C = np.array([[0.5], [100.0]])
    assert oracle.lloyd_iteration(X, C, empty="keep")["centroids"][1, 0] == 100.0
    assert np.isnan(oracle.lloyd_iteration(X, C, empty="nan")["centroids"][1, 0])


def test_near_tie_rule():
    d1 = np.array([1.0, 1.0, 100.0])
    d2 = np.array([1.000001, 1.1, 100.0005])
    assert oracle.near_tie_mask(d1, d2).tolist() == [True, False, True]
    assert oracle.label_mismatches([0, 1, 1], [1, 1, 0], d1, d2).tolist() == []
    assert oracle.label_mismatches([0, 0, 1], [1, 1, 0], d1, d2).tolist() == [1]


def test_golden_fixtures_reproduce(pkg, golden_index):
    for name in golden_index:
        z, meta = load_golden(name, pkg, golden_index)
        res = oracle.run_lloyd(z["X"], z["C0"], meta["iters"], n_shards=meta["shards"], form=meta["form"])
        for t, h in enumerate(res["history"]):
            assert oracle.rel_err(h["centroids"], z["centroids"][t]) < 1e-12, (name, t)
            assert np.array_equal(h["counts"], z["counts"][t]), (name, t)
            assert abs(h["inertia"] - z["inertia"][t]) <= 1e-10 * abs(z["inertia"][t]), (name, t)
            assert np.array_equal(h["labels"], z["labels"][t]), (name, t)
        assert np.array_equal(res["cluster_idx"], z["cluster_idx"]), name


def test_kmeanspp_oracle_is_seeded_and_spreads_centres(pkg):
    from oracle.kmeanspp_oracle import kmeanspp, splitmix64_uniforms

    u = splitmix64_uniforms(123128, 5)
    assert ((u >= 0) & (u < 1)).all() and len(set(u.tolist())) == 5
    X, centres = pkg.datagen.blobs_numpy(3000, 2, 6, seed=4)
    c1, rows1 = kmeanspp(X, 6, seed=9)
    c2, rows2 = kmeanspp(X, 6, seed=9)
    assert np.array_equal(rows1, rows2) and np.array_equal(c1, X[rows1])
    assert len(set(rows1.tolist())) == 6
    # D^2 sampling on well separated blobs picks (almost surely) one point of every blob
    owner = np.argmin(((c1[:, None, :] - centres[None, :, :]) ** 2).sum(2), axis=1)
    assert len(set(owner.tolist())) >= 5
    assert not np.array_equal(kmeanspp(X, 6, seed=10)[1], rows1)
