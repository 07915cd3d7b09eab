"""tensorflow-distributed-clustering_b200 -- the B200-native K-Means Lloyd hot path.

Only what the path needs: ``csrc/`` (sm_100a CUDA kernels + the C ABI of ``include/kmc.h``),
the ctypes binding, and the host-side mirror of the reference's Python entry point
(reference scripts/distribuitedClustering.py:180-318).  The directory name contains a hyphen,
so import it with ``importlib.import_module("tensorflow-distributed-clustering_b200")`` or
through the alias module ``tdc_b200`` at the repository root.
"""
from . import _lib, datagen, sharding  # noqa: F401
from .api import (  # noqa: F401
    LloydShard,
    allreduce_partials_single_process,
    distribuited_k_means,
    distribuited_k_means_spmd,
    kmeanspp_device,
    lloyd_loop,
    run_experiments,
)
from .sharding import parse_gpu_names, shard_ranges, shard_sizes  # noqa: F401

__all__ = [
    "LloydShard", "distribuited_k_means", "distribuited_k_means_spmd", "kmeanspp_device", "run_experiments", "lloyd_loop",
    "allreduce_partials_single_process", "shard_sizes", "shard_ranges", "parse_gpu_names",
    "datagen", "sharding",
]
