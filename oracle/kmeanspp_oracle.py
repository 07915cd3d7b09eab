"""CPU restatement of the seeded k-means++ (D^2 sampling) of csrc/kmeanspp.cu -- TEST INFRASTRUCTURE.

The reference calls sklearn's private, unseeded ``k_means_._init_centroids(data_batch, K,
init='k-means++')`` (scripts/distribuitedClustering.py:191); no reproducible output of it exists, so this
oracle pins the B200 routine instead: same splitmix64 stream, same two-level search (block sums of 1024
rows, then rows), float64 sums.  On data whose squared distances are exact in float32 (small integers)
the device must pick exactly the same rows.  PARITY UNPINNED against the reference itself.
"""
from __future__ import annotations

import numpy as np

BLOCK = 1024
MASK = (1 << 64) - 1


def splitmix64_uniforms(seed: int, k: int) -> np.ndarray:
    s = int(seed) & MASK
    out = np.empty(k, dtype=np.float64)
    for t in range(k):
        s = (s + 0x9E3779B97F4A7C15) & MASK
        z = s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK
        z = z ^ (z >> 31)
        out[t] = float(z >> 11) * (1.0 / 9007199254740992.0)
    return out


def kmeanspp(X: np.ndarray, K: int, seed: int):
    """Returns (centres float32 [K, D], chosen rows int64 [K])."""
    X = np.asarray(X, dtype=np.float32)
    n = X.shape[0]
    u = splitmix64_uniforms(seed, K)
    chosen = np.empty(K, dtype=np.int64)
    chosen[0] = min(int(u[0] * n), n - 1)
    mind = None
    n_blocks = (n + BLOCK - 1) // BLOCK
    per = (n_blocks + 1023) // 1024
    for t in range(1, K):
        diff = X.astype(np.float64) - X[chosen[t - 1]].astype(np.float64)
        d = (diff * diff).sum(1).astype(np.float32)          # exact for the integer test data
        mind = d if mind is None else np.minimum(mind, d)
        m64 = mind.astype(np.float64)
        block_sum = np.add.reduceat(m64, np.arange(0, n, BLOCK))
        slice_sum = np.add.reduceat(block_sum, np.arange(0, n_blocks, per))
        incl = np.cumsum(slice_sum)
        total = incl[-1]
        target = u[t] * total
        row = -1
        if total > 0.0:
            excl = incl - slice_sum
            hit = np.nonzero((incl >= target) & (excl < target))[0]
            if len(hit):
                i = int(hit[0])
                run, pick = excl[i], min(n_blocks, (i + 1) * per) - 1
                for b in range(i * per, min(n_blocks, (i + 1) * per)):
                    if run + block_sum[b] >= target:
                        pick = b
                        break
                    run += block_sum[b]
                blk, before = pick, run
            else:
                blk, before = n_blocks - 1, total - block_sum[-1]
        else:
            blk, before = 0, 0.0
        rows = m64[blk * BLOCK:(blk + 1) * BLOCK]
        incl_r = np.cumsum(rows)
        want = target - before
        hit = np.nonzero((rows > 0.0) & (incl_r >= want) & (incl_r - rows < want))[0]
        if len(hit):
            row = blk * BLOCK + int(hit[0])
        else:
            pos = np.nonzero(rows > 0.0)[0]
            row = blk * BLOCK + (int(pos[-1]) if len(pos) else 0)
        chosen[t] = row
    return X[chosen].copy(), chosen
