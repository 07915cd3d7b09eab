"""float64 NumPy restatement of the reference K-Means Lloyd path (TEST INFRASTRUCTURE).

Reference: ``/root/reference/scripts/distribuitedClustering.py`` (``distribuited_k_means``,
lines 180-294).  Every function cites the lines it restates.  The reference evaluates its
graph in float64 (it cannot run float32: line 248 multiplies a data-dtype tensor with a
float64 count); this oracle therefore evaluates the same math in float64 on whatever
values it is given (the float32 inputs of the product, widened exactly).

PARITY UNPINNED -- see ``oracle/__init__.py``.

Two update modes:

* ``lloyd_iteration``   direct per-cluster sums / counts.  This is the parity oracle.
* ``literal_iteration`` the reference's per-shard ``reduce_mean`` x ``bincount``
  reconstruction, including its NaN artefact for clusters that are empty on one shard
  (SURVEY.md section 8a note N2).  Used to document N2 and as the reference-shaped CPU timing.
"""
from __future__ import annotations

import numpy as np

__all__ = [
    "shard_sizes", "dist2", "assign", "lloyd_iteration", "literal_iteration",
    "run_lloyd", "near_tie_mask", "label_mismatches", "rel_err",
]


def shard_sizes(n_rows: int, n_shards: int) -> list[int]:
    """Row counts of ``np.array_split(data_batch, G)`` without materialising the shards.

    Reference: distribuitedClustering.py:184 (``sizes = [len(arg) for arg in
    np.array_split(data_batch, len(GPU_names))]``) and :197 (``tf.split(all_data, sizes, 0)``).
    The first ``N % G`` shards get ``N // G + 1`` rows, the rest ``N // G``.
    """
    if n_shards <= 0:
        raise ValueError("n_shards must be positive")
    base, extra = divmod(int(n_rows), int(n_shards))
    return [base + 1 if g < extra else base for g in range(n_shards)]


def dist2(X: np.ndarray, C: np.ndarray, form: str = "diff") -> np.ndarray:
    """Squared Euclidean distances ``[n, K]`` in float64.

    ``form='diff'`` restates distribuitedClustering.py:221-230 literally: broadcast both
    operands to ``[n, K, M]`` (the reference's tile+reshape is a broadcast), subtract,
    square, ``reduce_sum(axis=2)``.  ``form='gemm'`` is the float64 expansion
    ``|x|^2 - 2 x.c + |c|^2`` used only to make large oracle cases finish in seconds; in
    float64 it differs from 'diff' by ~1e-13 relative, far below every parity tolerance
    (checked in tests/test_oracle.py).
    """
    X = np.asarray(X, dtype=np.float64)
    C = np.asarray(C, dtype=np.float64)
    if form == "diff":
        sub = X[:, None, :] - C[None, :, :]          # :228 tf.subtract(rep_points, rep_centroids)
        return np.sum(sub * sub, axis=2)               # :229-230 square, reduce_sum(axis=2)
    if form == "gemm":
        xn = np.einsum("ij,ij->i", X, X)
        cn = np.einsum("ij,ij->i", C, C)
        d = xn[:, None] - 2.0 * (X @ C.T) + cn[None, :]
        return np.maximum(d, 0.0)
    raise ValueError(form)


def _chunk_rows(K: int, D: int, form: str) -> int:
    per_row = K * D * 8 * 2 if form == "diff" else K * 8 * 3
    return max(1, min(1 << 16, (96 << 20) // max(per_row, 1)))


def assign(X: np.ndarray, C: np.ndarray, form: str = "diff"):
    """Nearest-centroid labels plus best / second-best squared distances, chunked over rows.

    Reference: distribuitedClustering.py:234 ``best_centroids = tf.argmin(sum_squares, axis=1)``.
    Ties resolve to the lowest index (``np.argmin``; the Eigen reducer behind ``tf.argmin``
    does the same in practice).  NaN centroids never win a comparison: a NaN column is mapped
    to +inf before the argmin, which is what repeated ``<`` comparisons do.
    Returns ``labels int64 [n]``, ``d1 float64 [n]``, ``d2 float64 [n]`` (``d2 = inf`` if K == 1).
    """
    X = np.asarray(X)
    C = np.asarray(C, dtype=np.float64)
    n, K = X.shape[0], C.shape[0]
    labels = np.empty(n, dtype=np.int64)
    d1 = np.empty(n, dtype=np.float64)
    d2 = np.full(n, np.inf, dtype=np.float64)
    step = _chunk_rows(K, X.shape[1], form)
    for r0 in range(0, n, step):
        d = dist2(X[r0:r0 + step], C, form)
        d = np.where(np.isnan(d), np.inf, d)
        lab = np.argmin(d, axis=1)
        rows = np.arange(d.shape[0])
        labels[r0:r0 + step] = lab
        d1[r0:r0 + step] = d[rows, lab]
        if K > 1:
            d[rows, lab] = np.inf
            d2[r0:r0 + step] = d.min(axis=1)
    return labels, d1, d2


def lloyd_iteration(X, C, n_shards: int = 1, empty: str = "keep", form: str = "diff") -> dict:
    """One Lloyd iteration with direct sums/counts -- THE PARITY ORACLE.

    Mathematically what distribuitedClustering.py:221-263 computes wherever the reference
    is well defined (no cluster empty on any shard): ``c_new[k] = sum_{label_i=k} x_i /
    #{label_i=k}``.  Per-shard partial sums are formed exactly as rows are sharded at :184 and
    added shard by shard like ``tf.add_n`` at :257-258, so the float64 summation order follows
    the reference's.  ``empty='keep'`` leaves a globally empty cluster's centroid unchanged,
    ``empty='nan'`` reproduces the reference's 0/0 (:261).
    """
    X = np.asarray(X)
    C = np.asarray(C, dtype=np.float64)
    K, D = C.shape
    labels, d1, d2 = assign(X, C, form)
    sums = np.zeros((K, D), dtype=np.float64)
    counts = np.zeros(K, dtype=np.float64)
    r0 = 0
    for n_g in shard_sizes(X.shape[0], n_shards):
        lab_g = labels[r0:r0 + n_g]
        Xg = np.asarray(X[r0:r0 + n_g], dtype=np.float64)
        cnt_g = np.bincount(lab_g, minlength=K).astype(np.float64)      # :244-246
        sum_g = np.zeros((K, D), dtype=np.float64)
        if n_g:                                                         # = mean_g * cnt_g (:237-248)
            order = np.argsort(lab_g, kind="stable")                    # rows of one cluster, in row order
            starts = np.concatenate(([0], np.cumsum(cnt_g.astype(np.int64))[:-1]))
            live = cnt_g > 0
            sum_g[live] = np.add.reduceat(Xg[order], starts[live], axis=0)
        sums += sum_g                                                   # :258 add_n(partial_values)
        counts += cnt_g                                                 # :257 add_n(partial_directions)
        r0 += n_g
    with np.errstate(invalid="ignore", divide="ignore"):
        C_new = sums / counts[:, None]                                  # :260-261
    if empty == "keep":
        dead = counts == 0
        C_new[dead] = C[dead]
    elif empty != "nan":
        raise ValueError(empty)
    return {
        "centroids": C_new, "labels": labels, "counts": counts, "sums": sums,
        "inertia": float(d1.sum()), "d1": d1, "d2": d2,
        "shift2": float(np.nansum((C_new - C) ** 2)),
    }


def literal_iteration(X, C, n_shards: int = 1) -> dict:
    """One iteration exactly as the reference graph forms it, NaN artefacts included.

    distribuitedClustering.py:237-240 per shard and cluster: ``reduce_mean(gather(X, where(
    equal(best, c))))`` (mean of an empty set is NaN); :244-248 ``bincount`` and
    ``mean * count`` (NaN * 0 = NaN); :257-261 ``add_n`` over shards and the division.
    """
    X = np.asarray(X, dtype=np.float64)
    C = np.asarray(C, dtype=np.float64)
    K, D = C.shape
    labels = np.empty(X.shape[0], dtype=np.int64)
    sum_mu = np.zeros((D, K), dtype=np.float64)
    sum_direction = np.zeros(K, dtype=np.float64)
    r0 = 0
    for n_g in shard_sizes(X.shape[0], n_shards):
        Xg = X[r0:r0 + n_g]
        sum_squares = dist2(Xg, C, "diff")                              # :221-230
        best = np.argmin(np.where(np.isnan(sum_squares), np.inf, sum_squares), axis=1)  # :234
        labels[r0:r0 + n_g] = best
        means = []
        with np.errstate(invalid="ignore", divide="ignore"):
            for c in range(K):                                          # :237-240
                aux_points = Xg[np.where(best == c)[0]]
                means.append(aux_points.mean(axis=0) if len(aux_points) else np.full(D, np.nan))
            new_centroids = np.stack(means, 0)                          # :242
            y_count = np.bincount(best.astype(np.int32), minlength=K).astype(np.float64)  # :244-246
            partial_mu = new_centroids.T * y_count                      # :248
        sum_direction = sum_direction + y_count                         # :257
        sum_mu = sum_mu + partial_mu                                    # :258
        r0 += n_g
    with np.errstate(invalid="ignore", divide="ignore"):
        rep = np.tile(sum_direction, D).reshape(D, K)                   # :260
        new_centers = (sum_mu / rep).T                                  # :261
    return {"centroids": new_centers, "labels": labels, "counts": sum_direction}


def run_lloyd(X, C0, n_max_iters: int, n_shards: int = 1, empty: str = "keep",
              tol: float = 0.0, form: str = "diff", keep_labels: bool = True) -> dict:
    """The reference's iteration loop and result dictionary.

    distribuitedClustering.py:277-282: exactly ``n_max_iters`` updates (``tol`` is this
    project's addition; ``tol=0`` never stops early), then :255/:282 the label pass against
    the already-updated centroids.  Returns the reference's result keys (:284-292) plus a
    per-iteration ``history`` used by the parity tests: ``history[t]`` holds ``C_{t+1}``,
    the labels / counts / inertia computed against ``C_t`` (SURVEY.md section 8a note N4).
    """
    C = np.array(C0, dtype=np.float64, copy=True)
    history = []
    n_iter = 0
    for _ in range(int(n_max_iters)):
        it = lloyd_iteration(X, C, n_shards, empty, form)
        if not keep_labels:
            it = {k: v for k, v in it.items() if k not in ("labels", "d1", "d2")}
        history.append(it)
        C = it["centroids"]
        n_iter += 1
        if tol > 0.0 and it["shift2"] <= tol:
            break
    cluster_idx, d1, d2 = assign(X, C, form)                            # :255, :282
    last = shard_sizes(np.asarray(X).shape[0], n_shards)[-1]
    return {
        "end_center": C,
        "cluster_idx": cluster_idx,
        "centroids": history[-1]["labels"][-last:] if (history and keep_labels) else None,
        "init_center": np.asarray(C0),
        "n_iter": n_iter,
        "history": history,
        "final_d1": d1, "final_d2": d2,
        "inertia": history[-1]["inertia"] if history else float("nan"),
    }


def near_tie_mask(d1: np.ndarray, d2: np.ndarray, rtol: float = 1e-5) -> np.ndarray:
    """Rows whose two smallest oracle distances differ by <= ``rtol`` relative.

    BASELINE.json north_star: assignments must be identical "except on near-ties where
    |d1 - d2| <= 1e-5 relative"; relative to the larger of the two.
    """
    scale = np.maximum(np.abs(d1), np.abs(d2))
    return (d2 - d1) <= rtol * scale


def label_mismatches(labels_test, labels_oracle, d1, d2, rtol: float = 1e-5):
    """Indices where labels differ and the row is NOT a near-tie (must be empty for parity)."""
    bad = np.asarray(labels_test).astype(np.int64) != np.asarray(labels_oracle).astype(np.int64)
    return np.nonzero(bad & ~near_tie_mask(d1, d2, rtol))[0]


def rel_err(a, b) -> float:
    """max |a-b| / max(|b|) -- the 'within 1e-4 relative' measure used for centroids/inertia."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    denom = float(np.max(np.abs(b))) if b.size else 1.0
    return float(np.max(np.abs(a - b))) / (denom if denom > 0 else 1.0)
