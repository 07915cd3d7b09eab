"""Worker of tests/test_gpu_multi.py::test_torchrun_nccl_spmd_matches_oracle: one process per GPU (torchrun),
NCCL all-reduce of the partial buffers, result of rank 0 written as .npz for the parent test to check."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    out_path, n, d, k, iters = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    pkg = importlib.import_module("tensorflow-distributed-clustering_b200")
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    X, _ = pkg.datagen.blobs_numpy(n, d, k, seed=123128)
    C0 = X[:k].copy()
    r0, r1 = pkg.shard_ranges(n, world)[rank]
    res = pkg.distribuited_k_means_spmd(X[r0:r1], k, C0, iters, device=dev)
    gathered = [None] * world
    dist.all_gather_object(gathered, {"cluster_idx": res["cluster_idx"], "centroids": res["centroids"],
                                      "end_center": res["end_center"], "inertia": res["inertia"]})
    if rank == 0:
        np.savez(out_path, cluster_idx=np.concatenate([g["cluster_idx"] for g in gathered]),
                 last_centroids=gathered[-1]["centroids"], end_center=gathered[0]["end_center"],
                 end_centers_equal=all(np.array_equal(g["end_center"], gathered[0]["end_center"]) for g in gathered),
                 inertia=gathered[0]["inertia"])
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
