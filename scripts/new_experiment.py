#!/usr/bin/env python
"""Sweep driver: data size x K x GPU count, one `distribuitedClustering.py` process per cell.

Counterpart of the reference's scripts/new_experiment.py (:9-26 data generation with sklearn's
make_classification, :35-64 the nested size / K / GPU-count / method loop that shells out to the CLI under
nvprof and prints each cell's return code).  Differences, all forced by the platform: only
`distributedKMeans` exists here; the profiler is Nsight Compute (`--profile`, a launch list per cell under
`ncu_logs/`, read by tools/ncu_summary.py) instead of nvprof; the grid is given on the command line, the
reference's own grid being the default.

    python scripts/new_experiment.py                       # the reference's grid (25M..100M rows, D=5)
    python scripts/new_experiment.py --n_obs 1000000 --n_dim 2 --K 8 --gpus 1 --n_max_iters 20
"""
import argparse
import os
import subprocess
import sys
from datetime import datetime

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CLI = os.path.join(HERE, "distribuitedClustering.py")
DATA_SEED = 1826273          # reference :40
RUN_SEED = 123128            # reference :56


def make_data(filepath, n_obs, n_dim, seed):
    """Two-class make_classification data written as X / Y into an .npz (reference :9-26)."""
    from sklearn.datasets import make_classification

    if os.path.exists(filepath):
        os.remove(filepath)
    X, Y = make_classification(n_samples=n_obs, n_features=n_dim, n_informative=n_dim, n_redundant=0, n_classes=2,
                               n_clusters_per_class=1, shuffle=True, random_state=seed)
    np.savez(filepath, X=X, Y=Y)
    return True


def cell_command(n_obs, n_dim, K, n_gpus, n_max_iters, seed, log_csv, data_path, profile_log=None):
    cmd = [sys.executable, CLI, f"--n_obs={n_obs}", f"--n_dim={n_dim}", f"--K={K}", f"--n_GPUs={n_gpus}",
           f"--n_max_iters={n_max_iters}", f"--seed={seed}", f"--log_file={log_csv}",
           "--method_name=distributedKMeans", f"--data_file={data_path}"]
    if profile_log:
        cmd = ["ncu", "--metrics", "gpu__time_duration.sum", "--clock-control", "none", "--csv", "--log-file",
               profile_log] + cmd
    return cmd


def run_sweep(sizes, n_dim, Ks, gpu_counts, n_max_iters, log_csv, data_path, profile, runner=subprocess.call):
    """Returns [(cell name, return code)].  `runner(cmd) -> int` is injectable for tests."""
    results = []
    for n_obs in sizes:
        make_data(data_path, n_obs, n_dim, DATA_SEED)
        for K in Ks:
            for n_gpus in gpu_counts:
                name = f"distributedKMeans-GPUs{n_gpus}-n_obs{n_obs}-n_dims{n_dim}-K{K}.log"
                plog = None
                if profile:
                    os.makedirs("ncu_logs", exist_ok=True)
                    plog = os.path.join("ncu_logs", name)
                rc = runner(cell_command(n_obs, n_dim, K, n_gpus, n_max_iters, RUN_SEED, log_csv, data_path, plog))
                print(f"{name} - Return code: {rc}")
                results.append((name, rc))
    return results


if __name__ == "__main__":
    ap = argparse.ArgumentParser(description="size x K x GPU-count sweep of the B200 K-Means CLI")
    ap.add_argument("--n_obs", type=int, nargs="+", default=[100000000, 75000000, 50000000, 25000000])
    ap.add_argument("--n_dim", type=int, default=5)
    ap.add_argument("--K", type=int, nargs="+", default=[15, 12, 9, 6, 3])
    ap.add_argument("--gpus", type=int, nargs="+", default=[1, 2, 3, 4, 5, 6, 7, 8])
    ap.add_argument("--n_max_iters", type=int, default=20)
    ap.add_argument("--log_file", default="executions_log.csv")
    ap.add_argument("--data_file", default="class-data.npz")
    ap.add_argument("--profile", action="store_true", help="run every cell under ncu (launch list per cell)")
    a = ap.parse_args()
    t0 = datetime.now()
    run_sweep(a.n_obs, a.n_dim, a.K, a.gpus, a.n_max_iters, a.log_file, a.data_file, a.profile)
    print(datetime.now() - t0)
