"""float64 mirror of the oracle on the GPU -- TEST INFRASTRUCTURE, not product code.

SURVEY.md section 8c ("Oracle at full scale"): the NumPy oracle cannot evaluate N = 10^8 rows, so the
full-size checks use this torch-float64 restatement of the same arithmetic (reference
scripts/distribuitedClustering.py:221-234 distances + first-min argmin, :237-248 per-cluster sums and
counts), chunked over rows.  It is validated against `oracle.assign` / `oracle.lloyd_iteration` at small N
by tests/test_gpu_parity.py::test_gpu_mirror_matches_the_numpy_oracle.  Only tests/, smoke() and the
checking legs of bench.py may import it; none of the product's kernels are called here (plain torch ops).

Distances use the expanded form |x|^2 - 2 x.c + |c|^2 in float64 (one DGEMM per chunk): its rounding,
~1e-16 * (|x|^2 + |c|^2), is far below the 1e-5 near-tie window the parity rule leaves open.
"""
from __future__ import annotations

import torch


def assign_f64(X: torch.Tensor, C: torch.Tensor, chunk_rows: int = 1 << 18):
    """labels (int64), d1, d2 (float64: smallest and second-smallest squared distance) of every row of the
    float32/float64 CUDA tensor X [n, D] against C [K, D]; NaN centroids never win; lowest index on ties."""
    n = X.shape[0]
    C64 = C.to(torch.float64)
    cn = (C64 * C64).sum(1)
    labels = torch.empty(n, dtype=torch.int64, device=X.device)
    d1 = torch.empty(n, dtype=torch.float64, device=X.device)
    d2 = torch.empty(n, dtype=torch.float64, device=X.device)
    for r0 in range(0, n, chunk_rows):
        x = X[r0:r0 + chunk_rows].to(torch.float64)
        dist = (x * x).sum(1, keepdim=True) - 2.0 * (x @ C64.T) + cn[None, :]
        dist = torch.where(torch.isnan(dist), torch.full_like(dist, float("inf")), dist)
        if C64.shape[0] >= 2:
            two, idx = torch.topk(dist, 2, dim=1, largest=False, sorted=True)
            best = two[:, 0].clamp_min(0.0)
            # topk does not promise the lowest index among equal values: take the first column that attains it
            lab = (dist == two[:, :1]).to(torch.uint8).argmax(dim=1)
            d2[r0:r0 + chunk_rows] = two[:, 1].clamp_min(0.0)
        else:
            best = dist[:, 0].clamp_min(0.0)
            lab = torch.zeros(x.shape[0], dtype=torch.int64, device=X.device)
            d2[r0:r0 + chunk_rows] = float("inf")
        labels[r0:r0 + chunk_rows] = lab
        d1[r0:r0 + chunk_rows] = best
        del dist, x
    return labels, d1, d2


def lloyd_pass_f64(X: torch.Tensor, C: torch.Tensor, chunk_rows: int = 1 << 18):
    """One assignment + accumulation pass in float64: dict with labels, d1, d2, sums [K, D], counts [K], inertia."""
    K, D = C.shape
    labels, d1, d2 = assign_f64(X, C, chunk_rows)
    sums = torch.zeros((K, D), dtype=torch.float64, device=X.device)
    for r0 in range(0, X.shape[0], chunk_rows):
        sums.index_add_(0, labels[r0:r0 + chunk_rows], X[r0:r0 + chunk_rows].to(torch.float64))
    counts = torch.bincount(labels, minlength=K).to(torch.float64)
    return {"labels": labels, "d1": d1, "d2": d2, "sums": sums, "counts": counts, "inertia": d1.sum()}


def sums_for_labels_f64(X: torch.Tensor, labels: torch.Tensor, K: int, chunk_rows: int = 1 << 18) -> torch.Tensor:
    """Per-cluster float64 sums of the rows of X under GIVEN labels (reference :237-248 for those labels)."""
    sums = torch.zeros((K, X.shape[1]), dtype=torch.float64, device=X.device)
    for r0 in range(0, X.shape[0], chunk_rows):
        sums.index_add_(0, labels[r0:r0 + chunk_rows].to(torch.int64), X[r0:r0 + chunk_rows].to(torch.float64))
    return sums


def compare_pass(ref: dict, labels: torch.Tensor, sums: torch.Tensor, counts: torch.Tensor, inertia: float,
                 tie_rtol: float = 1e-5, X: torch.Tensor | None = None) -> dict:
    """The parity figures of one pass: labels outside near-ties, counts up to near-tie rows, sums and inertia.
    `sums_rel_err` compares with the oracle's sums, which differ by the rows the parity rule lets differ (labels
    inside near-ties); with `X` given, `sums_rel_err_own_labels` compares with the float64 sums of the SAME rows under
    the product's own labels, i.e. it checks the accumulation alone."""
    scale = torch.maximum(ref["d1"].abs(), ref["d2"].abs())
    near = (ref["d2"] - ref["d1"]) <= tie_rtol * scale
    diff = labels.to(torch.int64) != ref["labels"]
    n_near = int(near.sum().item())
    smax = float(ref["sums"].abs().max().item())
    extra = {}
    if X is not None:
        own = sums_for_labels_f64(X, labels, ref["sums"].shape[0])
        extra["sums_rel_err_own_labels"] = float((sums.to(torch.float64) - own).abs().max().item()) / (smax if smax > 0 else 1.0)
        del own
    return {
        **extra,
        "rows_checked": int(labels.numel()),
        "label_mismatches_outside_near_ties": int((diff & ~near).sum().item()),
        "label_differences_inside_near_ties": int((diff & near).sum().item()),
        "near_tie_rows": n_near,
        "counts_abs_diff": float((counts.to(torch.float64) - ref["counts"]).abs().sum().item()),
        "sums_rel_err": float((sums.to(torch.float64) - ref["sums"]).abs().max().item()) / (smax if smax > 0 else 1.0),
        "inertia_rel_err": abs(float(inertia) - float(ref["inertia"].item())) / max(float(ref["inertia"].item()), 1e-300),
    }
