"""CPU oracle for the data-parallel K-Means Lloyd path -- TEST INFRASTRUCTURE ONLY.

Nothing in the product path (the package ``tensorflow-distributed-clustering_b200``,
``scripts/distribuitedClustering.py``) may import this package.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` use it, and there only as the checker or the timed CPU baseline.

PARITY UNPINNED: the reference (``/root/reference/scripts/distribuitedClustering.py``)
is a TensorFlow 1.x graph; TensorFlow is not installed in this image and the reference
ships no tests, golden vectors or reproducible printed outputs for this path
(SURVEY.md section 4 and 8c).  The restatement in ``kmeans_oracle.py`` follows the
reference op for op (each function cites file:line) and is cross-checked against an
independent implementation of the same math, ``sklearn.cluster.KMeans(algorithm='lloyd')``
(``tests/golden/make_golden.py``), but it could not be checked against outputs of the
reference itself.

``gpu_mirror.py`` is the float64 restatement of the same arithmetic in torch on the GPU (SURVEY.md
section 8c, "Oracle at full scale"), validated against ``kmeans_oracle`` at small N by the GPU tests; it is
imported explicitly (``from oracle import gpu_mirror``) so that importing ``oracle`` needs no torch.
"""
from . import kmeanspp_oracle  # noqa: F401  (restatement of the device k-means++, csrc/kmeanspp.cu)
from .kmeans_oracle import (  # noqa: F401
    shard_sizes,
    dist2,
    assign,
    lloyd_iteration,
    literal_iteration,
    run_lloyd,
    near_tie_mask,
    label_mismatches,
    rel_err,
)
