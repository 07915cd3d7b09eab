#!/usr/bin/env python
"""Bring-up diagnostics for a GPU box (not a test, not a benchmark): runs one stage per process
so that a fault in one kernel cannot poison the CUDA context of the next stage.

    python tools/gpu_diag.py small|simt|tc_scores|tc_assign|accumulate|all
"""
import importlib
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _setup():
    import torch
    import oracle
    pkg = importlib.import_module("tensorflow-distributed-clustering_b200")
    dev = torch.device("cuda", 0)
    print("device:", torch.cuda.get_device_name(0), torch.cuda.get_device_capability(0), flush=True)
    return torch, oracle, pkg, dev


def _mk(pkg, torch, X, K, dev, path):
    s = pkg.LloydShard(X.shape[0], X.shape[1], K, dev, path=path)
    s.bind_points(torch.from_numpy(X).to(dev))
    return s


def stage_assign(path, shapes):
    torch, oracle, pkg, dev = _setup()
    for n, d, k in shapes:
        X, _ = pkg.datagen.blobs_numpy(n, d, max(2, k // 3), seed=n + d + k)
        C = X[:k].copy()
        s = _mk(pkg, torch, X, k, dev, path)
        s.set_centroids(C)
        t0 = time.time()
        lab = s.assign().cpu().numpy()
        ref, d1, d2 = oracle.assign(X, C, "gemm" if d * k > 20000 else "diff")
        bad = oracle.label_mismatches(lab, ref, d1, d2)
        st = s.stats()
        print(f"[{path}] n={n} d={d} k={k}: raw mismatches {(lab != ref).sum()}, outside near-ties {len(bad)}, "
              f"refined {st['ambiguous']}, {time.time() - t0:.2f}s", flush=True)
        if len(bad):
            print("   first bad rows", bad[:8], "got", lab[bad[:8]], "want", ref[bad[:8]], flush=True)
        s.accumulate()
        part = s.partials.cpu().numpy()
        it = oracle.lloyd_iteration(X, C.astype(np.float64), form="gemm" if d * k > 20000 else "diff")
        sums = part[:k * d].reshape(k, d)
        cnt = part[k * d:k * d + k]
        print(f"   accumulate: counts diff {np.abs(cnt - it['counts']).sum():.0f}, sums rel err "
              f"{oracle.rel_err(sums, it['sums']):.2e}, inertia rel err "
              f"{abs(part[k * d + k] - it['inertia']) / it['inertia']:.2e}", flush=True)
        s.update()
        print(f"   update: centroid rel err {oracle.rel_err(s.get_centroids(), it['centroids']):.2e}", flush=True)
        s.close()


def stage_tc_scores():
    torch, oracle, pkg, dev = _setup()
    for n, d, k in [(128, 64, 256), (256, 128, 256), (300, 128, 1024), (200, 256, 300), (130, 32, 100)]:
        X, _ = pkg.datagen.blobs_numpy(n, d, 16, seed=d)
        C = (X[:k].copy() if k <= n else np.concatenate([X] * (k // n + 1))[:k] * np.float32(0.9))
        s = _mk(pkg, torch, X, k, dev, "tc")
        s.set_centroids(C)
        got = s.debug_tc_scores().cpu().numpy().astype(np.float64)
        torch.cuda.synchronize()
        X64, C64 = X.astype(np.float64), C.astype(np.float64)
        want = (C64 * C64).sum(1)[None, :] - 2.0 * X64 @ C64.T
        bound = np.linalg.norm(X64, axis=1)[:, None] * (2 * np.linalg.norm(C64, axis=1))[None, :]
        err = np.abs(got - want) / bound
        ok = err < 1e-5
        print(f"[tc_scores] n={n} d={d} k={k}: max rel err {np.nanmax(err):.3e}, median {np.nanmedian(err):.3e}, "
              f"fraction ok {ok.mean():.4f}, nan {np.isnan(got).sum()}", flush=True)
        if ok.mean() < 1.0:
            np.set_printoptions(linewidth=200, precision=3, suppress=True)
            print("   got[:4,:8]\n", got[:4, :8], "\n   want[:4,:8]\n", want[:4, :8])
            print("   ok fraction by row block of 8 (first 16 blocks):", ok.reshape(-1, k)[: 128].reshape(16, 8, k).mean((1, 2)) if n >= 128 else "")
            print("   ok fraction by column block of 16 (first 16):", [round(float(ok[:, j:j + 16].mean()), 3) for j in range(0, min(k, 256), 16)])
            # is it a pure dot-product problem or a norm problem?
            dots_got = (got - (C64 * C64).sum(1)[None, :]) / -2.0
            dots_want = X64 @ C64.T
            print("   dot got/want ratio sample:", (dots_got[:2, :6] / dots_want[:2, :6]))
        s.close()


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "all"
    if mode == "small":
        stage_assign("small", [(5000, 2, 8), (4097, 3, 16), (3001, 8, 40), (1, 2, 1), (100000, 3, 16)])
    elif mode == "simt":
        stage_assign("simt", [(3000, 5, 7), (2500, 20, 100), (2049, 33, 65), (4096, 32, 256)])
    elif mode == "tc_scores":
        stage_tc_scores()
    elif mode == "tc_assign":
        stage_assign("tc", [(129, 64, 64), (4000, 16, 32), (3000, 32, 256), (2500, 64, 300), (20000, 128, 1024),
                            (1500, 100, 257), (1200, 160, 96), (1000, 256, 512)])
    elif mode == "all":
        rc = 0
        for m in ("small", "simt", "tc_scores", "tc_assign"):
            print(f"===== {m} =====", flush=True)
            r = subprocess.run([sys.executable, os.path.abspath(__file__), m], timeout=600)
            print(f"===== {m} exit {r.returncode} =====", flush=True)
            rc |= r.returncode
        sys.exit(rc)
    else:
        raise SystemExit("unknown mode")


if __name__ == "__main__":
    main()
