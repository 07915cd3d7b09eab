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


def _load_script(name):
    import importlib.util

    spec = importlib.util.spec_from_file_location(name.replace(".py", ""), os.path.join(ROOT, "scripts", name))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_cli_main_backs_off_on_out_of_memory_and_logs_a_csv_row(tmp_path):
    """Reference :331-360: a run that does not fit is retried with twice as many batches; the CSV row is
    appended with the timers of the run that went through."""
    dc = _load_script("distribuitedClustering.py")
    X = np.arange(64, dtype=np.float64).reshape(32, 2)
    data = tmp_path / "d.npz"
    np.savez(data, X=X, Y=np.zeros(32))
    seen = []

    def fake(batch, K, GPU_names, initial_centers, n_max_iters):
        seen.append(len(batch))
        if len(batch) > 8:
            raise RuntimeError("CUDA out of memory. Tried to allocate ...")
        assert np.array_equal(initial_centers, X[:3])            # :325 first K rows
        return {"end_center": np.zeros((K, 2)), "setup_time": 0.5, "initialization_time": 1.0, "computation_time": 2.0}

    log = tmp_path / "log.csv"
    rc = dc.main(32, 2, 3, ["/device:GPU:0"], 4, 99, str(log), "distributedKMeans", str(data), func=fake)
    assert rc == 0
    assert seen == [32, 16, 8, 8, 8, 8]                          # 1 batch fails, 2 batches fail at the first, 4 batches pass
    lines = log.read_text().strip().splitlines()
    assert lines[0].split(",") == dc.CSV_COLUMNS
    assert lines[1].split(",") == ["distributedKMeans", "99", "1", "3", "32", "2", "2.0", "4.0", "8.0", "4"]


def test_cli_main_writes_the_exception_name_and_returns_the_reference_status(tmp_path):
    """Reference :362-377: any other exception still appends a row with the exception's name in the timer
    columns; the exit status is 1 for a ValueError and 0 otherwise."""
    dc = _load_script("distribuitedClustering.py")
    data = tmp_path / "d.npz"
    np.savez(data, X=np.zeros((8, 2)), Y=np.zeros(8))
    log = tmp_path / "log.csv"

    def bad_value(*a):
        raise ValueError("K larger than N")

    def bad_other(*a):
        raise KeyError("x")

    assert dc.main(8, 2, 3, ["/device:GPU:0"], 5, 1, str(log), "distributedKMeans", str(data), func=bad_value) == 1
    assert dc.main(8, 2, 3, ["/device:GPU:0"], 5, 1, str(log), "distributedKMeans", str(data), func=bad_other) == 0
    assert dc.main(8, 2, 3, ["/device:GPU:0"], 5, 1, str(log), "distributedFuzzyCMeans", str(data), func=bad_other) == 0
    rows = [l.split(",") for l in log.read_text().strip().splitlines()[1:]]
    assert [r[6:9] for r in rows] == [["ValueError"] * 3, ["KeyError"] * 3, ["NotImplementedError"] * 3]
    assert all(r[9] == "5" for r in rows)
    assert dc.is_out_of_memory(MemoryError()) and not dc.is_out_of_memory(RuntimeError("illegal address"))


def test_sweep_walks_the_grid_and_reports_return_codes(tmp_path, monkeypatch):
    """Reference new_experiment.py:35-64: sizes x K x GPU counts, one CLI process per cell."""
    ne = _load_script("new_experiment.py")
    monkeypatch.chdir(tmp_path)
    cmds = []

    def runner(cmd):
        cmds.append(cmd)
        return len(cmds) % 2

    res = ne.run_sweep([200, 100], 3, [4, 2], [1, 2], 7, "log.csv", "data.npz", profile=False, runner=runner)
    assert len(res) == 8 and [rc for _, rc in res] == [1, 0] * 4
    assert res[0][0] == "distributedKMeans-GPUs1-n_obs200-n_dims3-K4.log"
    assert "--n_GPUs=2" in cmds[1] and "--K=4" in cmds[1] and "--seed=123128" in cmds[1] and "--n_max_iters=7" in cmds[1]
    d = np.load(tmp_path / "data.npz")
    assert d["X"].shape == (100, 3) and set(np.unique(d["Y"])) == {0, 1}       # the last size's data file
    prof = ne.cell_command(10, 2, 3, 1, 5, 1, "l.csv", "d.npz", "ncu_logs/x.log")
    assert prof[0] == "ncu" and "gpu__time_duration.sum" in prof
