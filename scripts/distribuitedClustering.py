#!/usr/bin/env python
"""Drop-in for the reference's scripts/distribuitedClustering.py -- K-Means part only.

Same module-level names and CLI flags as the reference (argparse block :411-491, `main` :320-409,
`run_experiments` :296-318, `distribuited_k_means` :180-294), but the work is done by the B200
library behind include/kmc.h instead of a TensorFlow graph.  Fuzzy C-Means (:72-178) is outside
this project's scope and is rejected.
"""
import argparse
import importlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
_pkg = importlib.import_module("tensorflow-distributed-clustering_b200")

distribuited_k_means = _pkg.distribuited_k_means
run_experiments = _pkg.run_experiments

CSV_COLUMNS = ["method_name", "seed", "num_GPUs", "K", "n_obs", "n_dim", "setup_time",
               "initialization_time", "computation_time", "n_iter"]


def get_available_gpus():
    """Reference :14-16 lists TensorFlow GPU device names; here the CUDA devices torch sees."""
    import torch

    return ["/device:GPU:%d" % i for i in range(torch.cuda.device_count())]


def main(n_obs, n_dim, K, GPU_names, n_max_iters, seed, log_file, method_name, data_file):
    """Reference :320-409: load X from the .npz, initial centres = X[0:K], run, append a CSV row."""
    if method_name != "distributedKMeans":
        raise SystemExit("only distributedKMeans is implemented on the B200 path")
    X = np.load(data_file)["X"]
    initial_centers = X[0:K, :]
    run_result = run_experiments([X], GPU_names, K, initial_centers, n_max_iters, distribuited_k_means)
    row = [method_name, seed, len(GPU_names), K, n_obs, n_dim, run_result["setup_time"],
           run_result["initialization_time"], run_result["computation_time"], run_result["n_iter"]]
    if not os.path.exists(log_file):
        with open(log_file, "w") as f:
            f.write(",".join(CSV_COLUMNS) + "\n")
    with open(log_file, "a") as f:
        f.write(",".join(str(v) for v in row) + "\n")
    print("log_file =", log_file)
    return 0


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
    a = ap.parse_args()
    names = get_available_gpus()
    if a.n_GPUs > len(names):
        ap.error("Number of GPUs Given is More Then Available")
    if a.n_GPUs <= 0:
        ap.error("Number of GPUs Given is Non Positive")
    sys.exit(main(a.n_obs, a.n_dim, a.K, names[:a.n_GPUs], a.n_max_iters, a.seed, a.log_file,
                  a.method_name, a.data_file))
