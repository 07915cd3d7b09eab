"""Row sharding of a batch across GPUs, following the reference's rule.

Reference: scripts/distribuitedClustering.py:184 takes the shard sizes from
``np.array_split(data_batch, len(GPU_names))`` and :197 applies them with ``tf.split``:
contiguous row ranges, the first ``N % G`` shards one row longer.  Nothing is copied here.
"""
from __future__ import annotations

import re


def shard_sizes(n_rows: int, n_shards: int) -> list[int]:
    if n_shards <= 0:
        raise ValueError("Number of GPUs Given is Non Positive")
    base, extra = divmod(int(n_rows), int(n_shards))
    return [base + 1 if g < extra else base for g in range(n_shards)]


def shard_ranges(n_rows: int, n_shards: int) -> list[tuple[int, int]]:
    out, r0 = [], 0
    for n_g in shard_sizes(n_rows, n_shards):
        out.append((r0, r0 + n_g))
        r0 += n_g
    return out


_DEV_RE = re.compile(r"(?:gpu|cuda)\s*:\s*(\d+)\s*$", re.IGNORECASE)


def parse_gpu_names(gpu_names, available: int) -> list[int]:
    """Device indices from the reference's ``GPU_names`` argument.

    Accepts what ``get_available_gpus()`` returns in the reference (:14-16: TensorFlow device
    strings such as '/device:GPU:0' or '/gpu:1'), torch-style 'cuda:0', plain ints, or a single
    int meaning "the first n GPUs".  Mirrors the checks of ``parse_valid_gpus_names`` (:58-70):
    more GPUs than available, or a non-positive count, is a ValueError.
    """
    if isinstance(gpu_names, (int,)) and not isinstance(gpu_names, bool):
        if gpu_names <= 0:
            raise ValueError("Number of GPUs Given is Non Positive")
        ids = list(range(gpu_names))
    else:
        ids = []
        for name in list(gpu_names):
            if isinstance(name, int):
                ids.append(int(name))
                continue
            if hasattr(name, "index") and hasattr(name, "type"):       # torch.device
                ids.append(int(name.index or 0))
                continue
            m = _DEV_RE.search(str(name))
            if not m:
                raise ValueError(f"cannot parse GPU name {name!r}")
            ids.append(int(m.group(1)))
        if not ids:
            raise ValueError("Number of GPUs Given is Non Positive")
    if len(set(ids)) != len(ids):
        raise ValueError("GPU names must be distinct")
    if len(ids) > available or any(i < 0 or i >= available for i in ids):
        raise ValueError("Number of GPUs Given is More Then Available")
    return ids
