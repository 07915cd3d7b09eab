"""Generate the golden fixtures under tests/golden/ (run once; outputs are committed).

The reference's TensorFlow graph cannot run in this image (no TensorFlow) and the
reference ships no golden vectors, so these fixtures pin the float64 oracle
(oracle/kmeans_oracle.py) instead, and every case is cross-checked here against an independent
implementation of the same mathematics, ``sklearn.cluster.KMeans(algorithm='lloyd')``.  Parity
with the reference itself stays UNPINNED (see oracle/__init__.py).

    python tests/golden/make_golden.py

Each ``<case>.npz`` holds the inputs (or their seed + SHA-256 when large) and, per iteration t,
the centroids C_{t+1}, counts, inertia and labels computed against C_t, the rows that are
near-ties (top-2 gap <= 1e-5 relative) and the final label pass.
"""
from __future__ import annotations

import hashlib
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

pkg = importlib.import_module("tensorflow-distributed-clustering_b200")
datagen = pkg.datagen
OUT = os.path.dirname(os.path.abspath(__file__))

CASES = [
    # name, kind, N, D, K, iters, shards, seed, form
    ("c1_blobs_d2_k8", "blobs", 4096, 2, 8, 6, 1, 123128, "diff"),
    ("c1_blobs_d2_k8_g3", "blobs", 4099, 2, 8, 6, 3, 800594, "diff"),
    ("c2_image_d3_k16", "image", 64 * 48, 3, 16, 6, 1, 123128, "diff"),
    ("c3_blobs_d32_k256", "blobs", 16384, 32, 256, 3, 2, 123128, "diff"),
    ("c4_blobs_d128_k1024", "blobs", 65536, 128, 1024, 2, 1, 123128, "gemm"),
    ("c5_blobs_d64_k4096", "blobs", 65536, 64, 4096, 1, 1, 800594, "gemm"),
    ("edge_k1_d1", "blobs", 257, 1, 1, 2, 2, 7, "diff"),
    ("edge_dup_centroids", "dup", 512, 4, 6, 3, 1, 11, "diff"),
]


def make_inputs(kind, N, D, K, seed):
    if kind == "blobs":
        X, _ = datagen.blobs_numpy(N, D, K, seed)
        return X, X[:K].copy()                        # reference convention: initial_centers = X[0:K]
    if kind == "image":
        X = datagen.image_numpy(48, 64, seed)
        return X, datagen.kmeanspp_init(X, K, seed)
    if kind == "dup":                                 # exact ties: centroid 1 duplicates centroid 0
        X, _ = datagen.blobs_numpy(N, D, K, seed)
        C0 = X[:K].copy()
        C0[1] = C0[0]
        return X, C0
    raise ValueError(kind)


def sklearn_check(X, C0, iters, res):
    from sklearn.cluster import KMeans

    km = KMeans(n_clusters=C0.shape[0], init=C0.astype(np.float64), n_init=1, algorithm="lloyd", tol=0.0,
                max_iter=iters).fit(X.astype(np.float64))
    err = oracle.rel_err(km.cluster_centers_, res["end_center"])
    mism = oracle.label_mismatches(km.labels_, res["cluster_idx"], res["final_d1"], res["final_d2"])
    return err, len(mism)


def main():
    index = {}
    for name, kind, N, D, K, iters, shards, seed, form in CASES:
        X, C0 = make_inputs(kind, N, D, K, seed)
        res = oracle.run_lloyd(X, C0, iters, n_shards=shards, empty="keep", form=form)
        hist = res["history"]
        any_empty = any((h["counts"] == 0).any() for h in hist)
        sk_err, sk_mism = (float("nan"), -1)
        if not any_empty and kind != "dup":
            sk_err, sk_mism = sklearn_check(X, C0, iters, res)
            assert sk_err < 1e-9, (name, sk_err)
            assert sk_mism == 0, (name, sk_mism)
        payload = {
            "C0": C0.astype(np.float32),
            "centroids": np.stack([h["centroids"] for h in hist]),
            "counts": np.stack([h["counts"] for h in hist]),
            "inertia": np.array([h["inertia"] for h in hist]),
            "labels": np.stack([h["labels"] for h in hist]).astype(np.int32),
            "near_tie": np.stack([oracle.near_tie_mask(h["d1"], h["d2"]) for h in hist]),
            "cluster_idx": res["cluster_idx"].astype(np.int32),
            "final_near_tie": oracle.near_tie_mask(res["final_d1"], res["final_d2"]),
        }
        store_x = X.nbytes <= 256 * 1024
        if store_x:
            payload["X"] = X
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **payload)
        index[name] = {
            "kind": kind, "N": int(X.shape[0]), "D": D, "K": K, "iters": iters, "shards": shards, "seed": seed,
            "form": form, "x_stored": store_x, "x_sha256": hashlib.sha256(X.tobytes()).hexdigest(),
            "any_empty": bool(any_empty), "sklearn_center_rel_err": sk_err, "sklearn_label_mismatches": sk_mism,
            "near_ties": int(sum(int(m.sum()) for m in payload["near_tie"])),
        }
        print(name, index[name])
    with open(os.path.join(OUT, "index.json"), "w") as f:
        json.dump(index, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
