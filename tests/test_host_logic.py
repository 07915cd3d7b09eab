"""CPU tests of the host side: the C-ABI library loads and exports every symbol that
include/kmc.h declares (no compute calls without a GPU), GPU-name parsing, argument handling of
the drop-in entry point, and loud failure when no CUDA device exists."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "kmc.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(kmc_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg._lib.load()
    declared = _declared_symbols()
    assert len(declared) >= 15
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/kmc.h but not exported"
    assert sorted(pkg._lib.SIGNATURES) == declared, "ctypes table and header disagree"
    assert lib.kmc_version() >= 1000
    assert lib.kmc_last_error() is not None


def test_create_rejects_bad_arguments_without_touching_a_gpu(pkg):
    lib = pkg._lib.load()
    ctx = ctypes.c_void_p()
    assert lib.kmc_create(ctypes.byref(ctx), 0, 10, 0, 3, 0) == -1          # D < 1
    assert b"D >= 1" in lib.kmc_last_error()
    assert lib.kmc_create(None, 0, 10, 2, 3, 0) == -1
    assert lib.kmc_workspace_bytes(None) == 0
    assert lib.kmc_partials_len(None) == 0


def test_parse_gpu_names(pkg):
    p = pkg.parse_gpu_names
    assert p(["/device:GPU:0", "/device:GPU:1"], 2) == [0, 1]
    assert p(["/gpu:1"], 2) == [1]
    assert p(["cuda:0"], 1) == [0]
    assert p(2, 8) == [0, 1]
    with pytest.raises(ValueError, match="More Then Available"):
        p(["/device:GPU:0", "/device:GPU:1"], 1)
    with pytest.raises(ValueError, match="Non Positive"):
        p(0, 4)
    with pytest.raises(ValueError):
        p([], 4)
    with pytest.raises(ValueError):
        p(["/cpu:0"], 4)
    with pytest.raises(ValueError):
        p(["/gpu:0", "/gpu:0"], 4)


def test_shard_ranges_are_contiguous(pkg):
    assert pkg.shard_ranges(10, 4) == [(0, 3), (3, 6), (6, 8), (8, 10)]


def test_entry_point_fails_loudly_without_cuda(pkg):
    import torch

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    X = np.zeros((16, 2), dtype=np.float32)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.distribuited_k_means(X, 2, ["/device:GPU:0"], X[:2], 3)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.LloydShard(16, 2, 2, "cuda:0")


def test_entry_point_validates_before_touching_the_device(pkg):
    X = np.zeros((4, 2), dtype=np.float32)
    with pytest.raises(ValueError, match="clusters but only"):
        pkg.distribuited_k_means(X, 9, 1, None, 3)
    with pytest.raises(ValueError):
        pkg.distribuited_k_means(np.zeros(4, dtype=np.float32), 2, 1, None, 3)


def test_run_experiments_averages_batch_centres(pkg):
    # reference :296-318 -- np.mean of the per-batch end centres, timers summed
    calls = []

    def fake(batch, K, GPU_names, initial_centers, n_max_iters):
        calls.append(len(batch))
        return {"end_center": np.full((K, 2), float(len(calls))), "setup_time": 1.0,
                "initialization_time": 2.0, "computation_time": 3.0}

    out = pkg.run_experiments([np.zeros((3, 2)), np.zeros((5, 2))], ["/gpu:0"], 2, np.zeros((2, 2)), 7, fake)
    assert calls == [3, 5]
    assert np.allclose(out["end_center"], 1.5)
    assert (out["setup_time"], out["initialization_time"], out["computation_time"], out["n_iter"]) == (2.0, 4.0, 6.0, 7)


def test_lloyd_loop_can_start_from_an_already_accumulated_first_iteration(pkg):
    """Streamed loading leaves the partial sums of iteration 1 in the shards: lloyd_loop(first_accumulated=True)
    must skip exactly that accumulate and otherwise be the plain loop (same centres, counts of iterations, tol)."""
    from test_distributed_gloo import OracleShard

    X, _ = pkg.datagen.blobs_numpy(4000, 3, 6, seed=4)
    C0 = X[:6].astype(np.float64)
    ranges = pkg.shard_ranges(len(X), 2)

    def reduce_on_cpu(shards):                       # the single-process default sums CUDA buffers with NCCL
        total = sum(s.partials for s in shards)
        for s in shards:
            s.partials.copy_(total)

    plain = [OracleShard(X[a:b], C0) for a, b in ranges]
    n_plain, inertia_plain, shift_plain = pkg.lloyd_loop(plain, 5, reduce_fn=reduce_on_cpu)
    pre = [OracleShard(X[a:b], C0) for a, b in ranges]
    calls = []
    for s in pre:
        s.accumulate()
        real = s.accumulate
        s.accumulate = (lambda real=real: (calls.append(1), real())[1])
    n_pre, inertia_pre, shift_pre = pkg.lloyd_loop(pre, 5, reduce_fn=reduce_on_cpu, first_accumulated=True)
    assert n_pre == n_plain == 5 and len(calls) == 2 * 4
    assert np.array_equal(pre[0].C, plain[0].C) and np.array_equal(pre[1].C, plain[1].C)
    assert inertia_pre == inertia_plain and shift_pre == shift_plain
    # tolerance stop right after the pre-accumulated iteration
    one = [OracleShard(X[a:b], C0) for a, b in ranges]
    for s in one:
        s.accumulate()
    n_one, _, shift_one = pkg.lloyd_loop(one, 50, tol=1e30, reduce_fn=reduce_on_cpu, first_accumulated=True)
    assert n_one == 1 and shift_one <= 1e30


def test_drop_in_script_exposes_reference_names():
    import importlib.util

    spec = importlib.util.spec_from_file_location("dc", os.path.join(ROOT, "scripts", "distribuitedClustering.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    for name in ("distribuited_k_means", "run_experiments", "main", "get_available_gpus"):
        assert callable(getattr(mod, name))
