"""Deterministic synthetic inputs for the BASELINE.json configurations.

Gaussian blobs (configs 1, 3, 4, 5): K_true centres drawn U(-spread, spread)^D from a seeded
generator, each row = a uniformly chosen centre + N(0, sigma^2), generated in float32.  Rows are
i.i.d., so ``X[0:K]`` (the reference's initial centres, scripts/distribuitedClustering.py:325)
is a random sample.  The reference's own generator is sklearn ``make_classification``
(notebooks/New-Distributed-KMeans.ipynb:56-73), which cannot produce K=8 classes in D=2.

Synthetic image (config 2): the colour-quantisation use of notebooks/Testing Images.ipynb:289-331
(``cv2.imread`` -> ``reshape(-1, 3)`` -> float): smooth colour gradients plus seeded noise,
quantised to integers 0..255, as a [H*W, 3] float array.
"""
from __future__ import annotations

import numpy as np


def blob_centers(k: int, d: int, seed: int, spread: float = 10.0) -> np.ndarray:
    rng = np.random.default_rng(seed)
    return rng.uniform(-spread, spread, size=(k, d)).astype(np.float32)


def blobs_numpy(n: int, d: int, k: int, seed: int, sigma: float = 1.0, spread: float = 10.0):
    """Host blobs: returns (X float32 [n, d], true_centers float32 [k, d])."""
    centers = blob_centers(k, d, seed, spread)
    rng = np.random.default_rng(seed + 1)
    which = rng.integers(0, k, size=n)
    X = centers[which] + rng.standard_normal(size=(n, d), dtype=np.float32) * np.float32(sigma)
    return np.ascontiguousarray(X, dtype=np.float32), centers


def blobs_torch(n: int, d: int, k: int, seed: int, device, sigma: float = 1.0, spread: float = 10.0,
                block_rows: int = 1 << 20, out=None, row0: int = 0, kind: str = "blobs"):
    """Rows [row0, row0 + n) of a device data set, generated straight into HBM (no host copy of the data).

    The data set is defined block by block: global block b (rows [b*block_rows, (b+1)*block_rows)) comes from a
    generator seeded with (seed, b), so any shard of it -- whatever the number of ranks -- holds the same rows
    as the corresponding slice of the single-GPU data set.  `kind`: "blobs" (a uniformly chosen centre +
    N(0, sigma^2)) or "uniform" (U(-spread, spread)^d, no cluster structure: the hard case for the tensor
    kernel's certification)."""
    import torch

    centers = torch.from_numpy(blob_centers(k, d, seed, spread)).to(device)
    X = out if out is not None else torch.empty((n, d), dtype=torch.float32, device=device)
    gen = torch.Generator(device=device)
    r_end = row0 + n
    tmp = None
    for b in range(row0 // block_rows, (r_end + block_rows - 1) // block_rows):
        g0, g1 = b * block_rows, (b + 1) * block_rows
        gen.manual_seed(int(seed) * 1000003 * 1009 + b)
        lo, hi = max(g0, row0), min(g1, r_end)
        whole = (lo == g0 and hi == g1)
        if whole:
            blk = X[lo - row0:hi - row0]
        else:
            if tmp is None:
                tmp = torch.empty((block_rows, d), dtype=torch.float32, device=device)
            blk = tmp
        if kind == "uniform":
            blk.uniform_(-spread, spread, generator=gen)
        else:
            which = torch.randint(0, k, (block_rows,), generator=gen, device=device)
            blk.normal_(0.0, sigma, generator=gen)
            blk += centers[which]
            del which
        if not whole:
            X[lo - row0:hi - row0].copy_(tmp[lo - g0:hi - g0])
    del tmp
    return X, centers


def image_numpy(height: int, width: int, seed: int) -> np.ndarray:
    """Synthetic RGB image as float32 pixels [height*width, 3] with integer values 0..255."""
    rng = np.random.default_rng(seed)
    yy = np.linspace(0.0, 1.0, height, dtype=np.float32)[:, None]
    xx = np.linspace(0.0, 1.0, width, dtype=np.float32)[None, :]
    r = 255.0 * (0.5 + 0.5 * np.sin(6.0 * xx + 2.0 * yy))
    g = 255.0 * (xx * 0.7 + yy * 0.3) * np.ones_like(r)
    b = 255.0 * (0.5 + 0.5 * np.cos(4.0 * yy - 3.0 * xx))
    img = np.stack([r, g, b], axis=-1).astype(np.float32)
    img += rng.normal(0.0, 12.0, size=img.shape).astype(np.float32)
    img = np.clip(np.rint(img), 0.0, 255.0)
    return np.ascontiguousarray(img.reshape(-1, 3), dtype=np.float32)


def image_torch(height: int, width: int, seed: int, device):
    import torch

    yy = torch.linspace(0.0, 1.0, height, device=device)[:, None]
    xx = torch.linspace(0.0, 1.0, width, device=device)[None, :]
    r = 255.0 * (0.5 + 0.5 * torch.sin(6.0 * xx + 2.0 * yy))
    g = 255.0 * (xx * 0.7 + yy * 0.3) * torch.ones_like(r)
    b = 255.0 * (0.5 + 0.5 * torch.cos(4.0 * yy - 3.0 * xx))
    img = torch.stack([r, g, b], dim=-1)
    gen = torch.Generator(device=device)
    gen.manual_seed(int(seed))
    img = img + torch.randn(img.shape, generator=gen, device=device) * 12.0
    return img.round_().clamp_(0.0, 255.0).reshape(-1, 3).contiguous()


def kmeanspp_init(X: np.ndarray, k: int, seed) -> np.ndarray:
    """Seeded k-means++ centres on the host.

    The reference HEAD calls sklearn's private ``k_means_._init_centroids(data, K,
    init='k-means++')`` unseeded (scripts/distribuitedClustering.py:191); the public successor is
    ``sklearn.cluster.kmeans_plusplus``.  Large inputs are subsampled (65 536 rows) first.
    """
    from sklearn.cluster import kmeans_plusplus

    X = np.asarray(X)
    rng = np.random.RandomState(None if seed is None else int(seed) % (2**32))
    if X.shape[0] > 65536:
        X = X[rng.choice(X.shape[0], 65536, replace=False)]
    centers, _ = kmeans_plusplus(np.asarray(X, dtype=np.float64), k, random_state=rng)
    return centers.astype(np.float32)
