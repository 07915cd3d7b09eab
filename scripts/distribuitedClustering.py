#!/usr/bin/env python
"""Drop-in for the reference's scripts/distribuitedClustering.py -- K-Means part only.

Same module-level names and CLI flags as the reference (argparse block :411-491, `main` :320-409,
`run_experiments` :296-318, `distribuited_k_means` :180-294), but the work is done by the B200
library behind include/kmc.h instead of a TensorFlow graph.  Fuzzy C-Means (:72-178) is outside
this project's scope and is rejected.

Behaviour kept from the reference's `main`:
  * initial centres are the first K rows of the data file (:325) -- `--init kmeans++` (an addition) draws
    them with the seeded on-device k-means++ instead, so that `--seed`, which the reference only logs, can
    select the start;
  * a run that does not fit in device memory is retried with twice as many batches (:331-360; the reference
    catches tf.errors.ResourceExhaustedError, here CUDA out-of-memory errors and KMC_E_NOMEM);
  * any other exception still appends a CSV row, with the exception's name in the timer columns, and the
    process exits 1 for a ValueError and 0 otherwise (:362-377) -- the sweep script relies on that code.
"""
import argparse
import importlib
import os
import sys
import traceback

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
_pkg = importlib.import_module("tensorflow-distributed-clustering_b200")

distribuited_k_means = _pkg.distribuited_k_means
run_experiments = _pkg.run_experiments

CSV_COLUMNS = ["method_name", "seed", "num_GPUs", "K", "n_obs", "n_dim", "setup_time",
               "initialization_time", "computation_time", "n_iter"]
MAX_BATCHES = 1 << 16


def get_available_gpus():
    """Reference :14-16 lists TensorFlow GPU device names; here the CUDA devices torch sees."""
    import torch

    return ["/device:GPU:%d" % i for i in range(torch.cuda.device_count())]


def is_out_of_memory(exc) -> bool:
    """CUDA / library out-of-memory conditions: the analogue of tf.errors.ResourceExhaustedError (:357)."""
    try:
        import torch
        if isinstance(exc, torch.cuda.OutOfMemoryError):
            return True
    except Exception:
        pass
    if isinstance(exc, MemoryError):
        return True
    text = str(exc).lower()
    return isinstance(exc, RuntimeError) and ("out of memory" in text or "kmc_e_nomem" in text or "cudaerrormemoryallocation" in text)


def append_csv_row(log_file, row):
    if not os.path.exists(log_file) or os.path.getsize(log_file) == 0:
        with open(log_file, "w") as f:
            f.write(",".join(CSV_COLUMNS) + "\n")
    with open(log_file, "a") as f:
        f.write(",".join(str(v) for v in row) + "\n")


def main(n_obs, n_dim, K, GPU_names, n_max_iters, seed, log_file, method_name, data_file, init="first",
         func=None):
    """Reference :320-409: load X from the .npz, choose the initial centres, run (halving the batch size while
    the run does not fit), append one CSV row whatever happened, return the exit status."""
    func = func or distribuited_k_means
    return_status = 0
    num_batches = 1
    run_result = None
    try:
        if method_name != "distributedKMeans":
            raise NotImplementedError("only distributedKMeans is implemented on the B200 path")
        X = np.load(data_file)["X"]
        if init == "kmeans++":
            initial_centers, _ = _pkg.kmeanspp_device(X, K, seed)
        else:
            initial_centers = X[0:K, :]
        while run_result is None:
            batches = np.array_split(X, num_batches)                              # :334
            print(num_batches)
            try:
                run_result = run_experiments(batches, GPU_names, K, initial_centers, n_max_iters, func)
            except Exception as e:                                                # noqa: BLE001
                if not is_out_of_memory(e) or num_batches >= MAX_BATCHES:
                    raise
                print("caught out-of-memory error")
                num_batches *= 2                                                  # :359
    except Exception:                                                             # noqa: BLE001  (:362-377)
        exc_type, _exc, exc_tb = sys.exc_info()
        traceback.print_tb(exc_tb)
        name = exc_type.__name__
        run_result = {"setup_time": name, "initialization_time": name, "computation_time": name, "n_iter": n_max_iters}
        return_status = 1 if name == "ValueError" else 0
    append_csv_row(log_file, [method_name, seed, len(GPU_names), K, n_obs, n_dim, run_result["setup_time"],
                              run_result["initialization_time"], run_result["computation_time"], run_result["n_iter"]])
    print("log_file =", log_file)
    return return_status


if __name__ == "__main__":
    ap = argparse.ArgumentParser(description="K-Means (Lloyd) on B200 GPUs")
    ap.add_argument("--n_obs", type=int, required=True)
    ap.add_argument("--n_dim", type=int, required=True)
    ap.add_argument("--K", type=int, required=True)
    ap.add_argument("--n_GPUs", type=int, required=True)
    ap.add_argument("--n_max_iters", type=int, required=True)
    ap.add_argument("--seed", type=int, required=True)
    ap.add_argument("--log_file", type=str, required=True)
    ap.add_argument("--method_name", type=str, required=True)
    ap.add_argument("--data_file", type=str, required=True)
    ap.add_argument("--init", choices=["first", "kmeans++"], default="first",
                    help="first: X[0:K] like the reference; kmeans++: seeded k-means++ on the GPU (uses --seed)")
    a = ap.parse_args()
    names = get_available_gpus()
    if a.n_GPUs > len(names):
        ap.error("Number of GPUs Given is More Then Available")
    if a.n_GPUs <= 0:
        ap.error("Number of GPUs Given is Non Positive")
    if not os.path.exists(a.data_file):
        ap.error("Unable to find data file !!!")
    sys.exit(main(a.n_obs, a.n_dim, a.K, names[:a.n_GPUs], a.n_max_iters, a.seed, a.log_file,
                  a.method_name, a.data_file, a.init))
