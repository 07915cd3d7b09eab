"""Host side of the B200 K-Means Lloyd path: thin Python over the C ABI of libkmc_b200.so.

Mirrors the reference's entry point ``distribuited_k_means`` and its caller ``run_experiments``
(reference scripts/distribuitedClustering.py:180-294 and :296-318): same names, positional forms,
result-dictionary keys and timer meanings.  PyTorch is used only for device memory, streams and
the cross-GPU all-reduce; every arithmetic step runs in hand-written sm_100a kernels behind
``include/kmc.h``.  There is no CPU or PyTorch fallback: without the built library or without a
B200 these functions raise.
"""
from __future__ import annotations

import ctypes
import time
from ctypes import byref, c_double, c_int, c_void_p

import numpy as np

from . import _lib as L
from .sharding import parse_gpu_names, shard_ranges


def _torch():
    import torch
    return torch


class LloydShard:
    """One contiguous shard of rows resident on one GPU, driven through one ``kmc_ctx``.

    The reference's per-GPU scope (``with tf.device(GPU_name)``, :201-251) holds the same state:
    the shard ``X = tf.Variable(X_mat)`` and the ops that produce its partial sums and counts.
    """

    def __init__(self, n_rows: int, D: int, K: int, device, empty: str = "keep", path: str = "auto"):
        torch = _torch()
        self.lib = L.load()
        if not torch.cuda.is_available():
            raise RuntimeError("libkmc_b200 needs a CUDA device (sm_100a); no CPU fallback exists")
        self.device = torch.device(device)
        self.n, self.D, self.K = int(n_rows), int(D), int(K)
        flags = L.PATH_FLAGS[path] | (L.KMC_EMPTY_NAN if empty == "nan" else L.KMC_EMPTY_KEEP)
        if empty not in ("keep", "nan"):
            raise ValueError("empty must be 'keep' or 'nan'")
        self._ctx = c_void_p()
        L.check(self.lib.kmc_create(byref(self._ctx), self.device.index or 0, self.n, self.D, self.K, flags))
        nbytes = int(self.lib.kmc_workspace_bytes(self._ctx))
        with torch.cuda.device(self.device):
            self.workspace = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=self.device)
            L.check(self.lib.kmc_bind_workspace(self._ctx, c_void_p(self.workspace.data_ptr()), nbytes))
        self.points = None
        plen = int(self.lib.kmc_partials_len(self._ctx))
        off = int(self.lib.kmc_partials_ptr(self._ctx)) - self.workspace.data_ptr()
        # [K*D sums | K counts | inertia | reserved] as float64, living inside the workspace
        self.partials = self.workspace[off:off + plen * 8].view(torch.float64)
        coff = int(self.lib.kmc_centroids_ptr(self._ctx)) - self.workspace.data_ptr()
        self.centroids_dev = self.workspace[coff:coff + self.K * self.D * 4].view(torch.float32).view(self.K, self.D)

    # -- lifetime -------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            self.lib.kmc_destroy(self._ctx)
            self._ctx = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return c_void_p(_torch().cuda.current_stream(self.device).cuda_stream)

    # -- data -----------------------------------------------------------------------------------
    def bind_points(self, points):
        """Borrow a float32 CUDA tensor [n_rows, D] (kept alive by this object)."""
        torch = _torch()
        if points.device != self.device or points.dtype != torch.float32 or not points.is_contiguous():
            raise ValueError("points must be a contiguous float32 tensor on the shard's device")
        if tuple(points.shape) != (self.n, self.D):
            raise ValueError(f"points shape {tuple(points.shape)} != {(self.n, self.D)}")
        self.points = points
        with torch.cuda.device(self.device):
            L.check(self.lib.kmc_bind_points(self._ctx, c_void_p(points.data_ptr()), self._stream()))

    def stage_from_host(self, host, row0: int, batch_size=None, out=None):
        """Pinned, chunked H2D of rows [row0, row0 + n_rows) of a host array, then bind them.

        Replaces feeding ``all_data`` and the per-GPU variable initialisation (:196-197, :217, :273).
        ``host`` is a C-contiguous float32/float64 NumPy array or CPU tensor of shape [N, D].
        """
        torch = _torch()
        if isinstance(host, torch.Tensor):
            arr_ptr, dt = host.data_ptr(), host.dtype
            is64 = dt == torch.float64
            if dt not in (torch.float32, torch.float64) or not host.is_contiguous():
                raise ValueError("host tensor must be contiguous float32/float64")
        else:
            if host.dtype not in (np.float32, np.float64) or not host.flags.c_contiguous:
                raise ValueError("host array must be C-contiguous float32/float64")
            arr_ptr, is64 = host.ctypes.data, host.dtype == np.float64
        with torch.cuda.device(self.device):
            if out is not None:
                pts = out[:self.n]
            else:
                pts = torch.empty((self.n, self.D), dtype=torch.float32, device=self.device)
            step = self.n if not batch_size else max(1, int(batch_size))
            for r in range(0, self.n, max(step, 1)):
                nr = min(step, self.n - r)
                dst = c_void_p(pts.data_ptr() + r * self.D * 4)
                L.check(self.lib.kmc_stage_points(self._ctx, dst, c_void_p(arr_ptr),
                                                  L.KMC_DTYPE_F64 if is64 else L.KMC_DTYPE_F32,
                                                  row0 + r, nr, self._stream()))
        self.bind_points(pts)

    def stage_and_accumulate_from_host(self, host, row0: int, piece_rows=None) -> bool:
        """H2D of rows [row0, row0 + n_rows) in pieces on a copy stream, each piece assigned and accumulated
        (the first Lloyd iteration; centroids must be set) while the next piece is on its way.  Leaves the
        shard as after stage_from_host() + accumulate().  Returns False -- having done nothing -- where streamed
        loading does not apply (small-D / SIMT path); falls back to a plain rebind + accumulate() if the plane
        scale fixed by the first piece turned out too small for later rows."""
        torch = _torch()
        if self.stats()["path"] != "tc" or self.n < (1 << 16):
            return False
        if isinstance(host, torch.Tensor):
            arr_ptr, is64 = host.data_ptr(), host.dtype == torch.float64
            if host.dtype not in (torch.float32, torch.float64) or not host.is_contiguous():
                raise ValueError("host tensor must be contiguous float32/float64")
        else:
            if host.dtype not in (np.float32, np.float64) or not host.flags.c_contiguous:
                raise ValueError("host array must be C-contiguous float32/float64")
            arr_ptr, is64 = host.ctypes.data, host.dtype == np.float64
        piece = int(piece_rows) if piece_rows else max(1 << 20, -(-self.n // 8))
        piece = -(-piece // 128) * 128
        with torch.cuda.device(self.device):
            pts = torch.empty((self.n, self.D), dtype=torch.float32, device=self.device)
            compute = torch.cuda.current_stream(self.device)
            copy = torch.cuda.Stream(self.device)
            copy.wait_stream(compute)
            L.check(self.lib.kmc_bind_points_streamed(self._ctx, c_void_p(pts.data_ptr()), self._stream()))
            self.points = pts
            for r0 in range(0, self.n, piece):
                r1 = min(self.n, r0 + piece)
                L.check(self.lib.kmc_stage_points(self._ctx, c_void_p(pts.data_ptr() + r0 * self.D * 4), c_void_p(arr_ptr),
                                                  L.KMC_DTYPE_F64 if is64 else L.KMC_DTYPE_F32, row0 + r0, r1 - r0,
                                                  c_void_p(copy.cuda_stream)))
                landed = torch.cuda.Event()
                landed.record(copy)
                compute.wait_event(landed)
                L.check(self.lib.kmc_accumulate_rows(self._ctx, r0, r1, self._stream()))
            if self.stats()["x_scaled_max"] >= 32768.0:      # later rows overflowed the planes' fp16 range
                self.bind_points(pts)
                self.accumulate()
        return True

    def set_centroids(self, C):
        C = np.ascontiguousarray(np.asarray(C, dtype=np.float32))
        if C.shape != (self.K, self.D):
            raise ValueError(f"initial centres shape {C.shape} != {(self.K, self.D)}")
        with _torch().cuda.device(self.device):
            L.check(self.lib.kmc_set_centroids(self._ctx, c_void_p(C.ctypes.data), self._stream()))

    def get_centroids(self) -> np.ndarray:
        out = np.empty((self.K, self.D), dtype=np.float32)
        with _torch().cuda.device(self.device):
            L.check(self.lib.kmc_get_centroids(self._ctx, c_void_p(out.ctypes.data), self._stream()))
        return out

    # -- iteration ------------------------------------------------------------------------------
    def accumulate(self):
        with _torch().cuda.device(self.device):
            L.check(self.lib.kmc_accumulate(self._ctx, self._stream()))

    def accumulate_add(self):
        """Like accumulate(), but adds this batch to the partial buffer (out-of-core batches)."""
        with _torch().cuda.device(self.device):
            L.check(self.lib.kmc_accumulate_add(self._ctx, self._stream()))

    def resize(self, n_rows: int):
        """Rows of the next batch (<= the size the shard was created for); bind its points afterwards."""
        L.check(self.lib.kmc_resize(self._ctx, int(n_rows)))
        self.n = int(n_rows)
        self.points = None

    def update(self):
        with _torch().cuda.device(self.device):
            L.check(self.lib.kmc_update(self._ctx, self._stream()))

    def iterate(self, max_iters: int, tol: float = 0.0):
        n_iter, inertia, shift2 = c_int(0), c_double(0.0), c_double(0.0)
        with _torch().cuda.device(self.device):
            L.check(self.lib.kmc_iterate(self._ctx, int(max_iters), float(tol), byref(n_iter), byref(inertia),
                                         byref(shift2), self._stream()))
        return n_iter.value, inertia.value, shift2.value

    def assign(self, out=None):
        torch = _torch()
        with torch.cuda.device(self.device):
            if out is None:
                out = torch.empty(self.n, dtype=torch.int32, device=self.device)
            L.check(self.lib.kmc_assign(self._ctx, c_void_p(out.data_ptr()), self._stream()))
        return out

    def last_labels(self, out=None):
        """int32 CUDA tensor [n_rows]: the labels of the last accumulate() pass, i.e. the argmin against the
        centroids BEFORE the update that followed it -- what the reference fetches as `best_centroids` in the
        same sess.run as the update (:279) and returns as `centroids` (:286)."""
        torch = _torch()
        with torch.cuda.device(self.device):
            if out is None:
                out = torch.empty(self.n, dtype=torch.int32, device=self.device)
            L.check(self.lib.kmc_last_labels(self._ctx, c_void_p(out.data_ptr()), self._stream()))
        return out

    def recolour(self, dev_labels=None, out=None):
        """Every row replaced by its centroid (`center[label]`, the reference notebook's colour quantisation
        output): float32 CUDA tensor [n_rows, D].  `dev_labels`: int32 labels from assign() (computed if None)."""
        torch = _torch()
        with torch.cuda.device(self.device):
            if dev_labels is None:
                dev_labels = self.assign()
            if out is None:
                out = torch.empty((self.n, self.D), dtype=torch.float32, device=self.device)
            L.check(self.lib.kmc_gather_centroids(self._ctx, c_void_p(dev_labels.data_ptr()), c_void_p(out.data_ptr()),
                                                  self._stream()))
        return out

    def labels_to_host(self, dev_labels, out: np.ndarray) -> np.ndarray:
        """Copy int32 device labels (from assign()) into the host array `out` (int32 or int64, contiguous):
        widened on the device, staged through pinned buffers, no pinned allocation of the result."""
        if out.dtype not in (np.int32, np.int64) or not out.flags.c_contiguous or out.shape != (dev_labels.numel(),):
            raise ValueError("out must be a contiguous int32/int64 array with one entry per label")
        with _torch().cuda.device(self.device):
            L.check(self.lib.kmc_labels_to_host(self._ctx, c_void_p(dev_labels.data_ptr()), c_void_p(out.ctypes.data),
                                                int(out.dtype.itemsize), int(out.shape[0]), self._stream()))
        return out

    def stats(self) -> dict:
        buf = (c_double * L.KMC_STAT_COUNT)()
        with _torch().cuda.device(self.device):
            L.check(self.lib.kmc_read_stats(self._ctx, buf, self._stream()))
        return {
            "inertia": buf[L.KMC_STAT_INERTIA], "shift2": buf[L.KMC_STAT_SHIFT2],
            "ambiguous": int(buf[L.KMC_STAT_AMBIGUOUS]), "x_scaled_max": buf[L.KMC_STAT_X_SCALED_MAX],
            "path": L.PATH_NAMES.get(int(buf[L.KMC_STAT_KERNEL_PATH]), "?"),
            "launches": int(buf[L.KMC_STAT_LAUNCHES]),
            "changed_rows": int(buf[L.KMC_STAT_CHANGED]), "full_pass": bool(buf[L.KMC_STAT_FULL_PASS]),
            "walk_overflow": int(buf[L.KMC_STAT_WALK_OVERFLOW]),
        }

    def set_param(self, param: int, value: float):
        L.check(self.lib.kmc_set_param(self._ctx, int(param), float(value)))

    def enable_profile(self, on: bool = True):
        self.set_param(L.KMC_PARAM_PROFILE, 1.0 if on else 0.0)

    def phase_ms(self) -> dict:
        """Average device milliseconds per phase since the last call (needs enable_profile)."""
        from ctypes import c_float
        buf = (c_float * L.KMC_PHASE_COUNT)()
        L.check(self.lib.kmc_get_phase_ms(self._ctx, buf, L.KMC_PHASE_COUNT))
        return {"assign": buf[0], "refine": buf[1], "accumulate": buf[2], "update": buf[3], "calls": int(buf[4])}

    def debug_tc_scores(self):
        torch = _torch()
        with torch.cuda.device(self.device):
            out = torch.empty((self.n, self.K), dtype=torch.float32, device=self.device)
            L.check(self.lib.kmc_debug_tc_scores(self._ctx, c_void_p(out.data_ptr()), self._stream()))
        return out


def kmeanspp_device(points, K: int, seed, max_rows: int = 1 << 20):
    """Seeded k-means++ on the GPU (kmc_kmeanspp).  `points`: [N, D] float CUDA tensor, or a host array /
    CPU tensor, in which case a seeded uniform subsample of at most `max_rows` rows is uploaded first.
    Returns (centres float32 ndarray [K, D], chosen row indices into `points`).
    Replaces the reference's host-side, unseeded `k_means_._init_centroids(..., init='k-means++')` (:191)."""
    torch = _torch()
    lib = L.load()
    seed = 0 if seed is None else int(seed) & ((1 << 64) - 1)
    N = int(points.shape[0])
    rng = np.random.default_rng(seed)
    sub_idx = None
    if isinstance(points, torch.Tensor) and points.is_cuda:
        device = points.device
        if N > max_rows:
            sub_idx = np.sort(rng.choice(N, max_rows, replace=False))
            pts = points[torch.from_numpy(sub_idx).to(device)].to(torch.float32).contiguous()
        else:
            pts = points.to(torch.float32).contiguous()
    else:
        device = torch.device("cuda", torch.cuda.current_device())
        host = points.numpy() if isinstance(points, torch.Tensor) else np.asarray(points)
        if N > max_rows:
            sub_idx = np.sort(rng.choice(N, max_rows, replace=False))
            host = host[sub_idx]
        pts = torch.from_numpy(np.ascontiguousarray(host, dtype=np.float32)).to(device)
    n, D = int(pts.shape[0]), int(pts.shape[1])
    if K > n:
        raise ValueError(f"K={K} clusters but only {n} candidate rows")
    with torch.cuda.device(device):
        nbytes = int(lib.kmc_kmeanspp_scratch_bytes(n, int(K)))
        scratch = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)
        cent = torch.empty((int(K), D), dtype=torch.float32, device=device)
        chosen = np.empty(int(K), dtype=np.int64)
        stream = c_void_p(torch.cuda.current_stream(device).cuda_stream)
        L.check(lib.kmc_kmeanspp(device.index or 0, c_void_p(pts.data_ptr()), n, D, int(K), ctypes.c_uint64(seed),
                                 c_void_p(cent.data_ptr()), c_void_p(chosen.ctypes.data), c_void_p(scratch.data_ptr()),
                                 nbytes, stream))
        centres = cent.cpu().numpy()
    if sub_idx is not None:
        chosen = sub_idx[chosen]
    return centres, chosen


# ---------------------------------------------------------------------------------------------
# Cross-shard reduction (the reference's add_n on /cpu:0, :257-258)
# ---------------------------------------------------------------------------------------------
def allreduce_partials_single_process(shards):
    """Sum the partial buffers of several shards driven by ONE process (one ctx per GPU)."""
    if len(shards) == 1:
        return
    torch = _torch()
    torch.cuda.nccl.all_reduce([s.partials for s in shards])


def lloyd_loop(shards, n_max_iters: int, tol: float = 0.0, reduce_fn=None, first_accumulated: bool = False):
    """Iterate ``n_max_iters`` times (reference :277-280): accumulate on every shard, sum the
    partial buffers across shards, update on every shard (identical arithmetic on identical
    inputs keeps the replicated centroids bit-identical, so no broadcast is needed).

    ``shards`` is a list of shard engines in this process (objects with ``accumulate()``,
    ``update()``, ``partials`` and ``stats()``); ``reduce_fn(shards)`` sums their ``partials`` in
    place -- across the list, and across processes when every rank runs this loop on its own
    shard (``reduce_fn = lambda s: torch.distributed.all_reduce(s[0].partials)``).
    ``tol > 0`` stops when the squared centre shift drops to ``tol`` (same decision on every
    rank: the shift is computed from the reduced buffer, which NCCL leaves bit-identical on all ranks, in a fixed
    summation order, see update_kernel).  ``first_accumulated``: the shards already hold
    the partial sums of the first iteration (streamed loading).  Returns (iterations, inertia, shift2).
    """
    if reduce_fn is None:
        reduce_fn = allreduce_partials_single_process
    n_iter, inertia, shift2 = 0, float("nan"), float("nan")
    for it in range(int(n_max_iters)):
        if not (first_accumulated and it == 0):
            for s in shards:
                s.accumulate()
        reduce_fn(shards)
        for s in shards:
            s.update()
        n_iter += 1
        if tol > 0.0:
            st = shards[0].stats()
            inertia, shift2 = st["inertia"], st["shift2"]
            if shift2 <= tol:
                break
    if shards:
        st = shards[0].stats()
        inertia, shift2 = st["inertia"], st["shift2"]
    return n_iter, inertia, shift2


# ---------------------------------------------------------------------------------------------
# Drop-in entry points
# ---------------------------------------------------------------------------------------------
def _as_host_matrix(data_batch):
    torch = _torch()
    if isinstance(data_batch, torch.Tensor):
        if data_batch.dim() != 2:
            raise ValueError("data_batch must be [N, D]")
        return data_batch
    arr = np.asarray(data_batch)
    if arr.ndim != 2:
        raise ValueError("data_batch must be [N, D]")
    if arr.dtype not in (np.float32, np.float64):
        arr = arr.astype(np.float64)
    return np.ascontiguousarray(arr)


def _prefaulted_labels(n: int):
    """The int64 label array the call returns, with its pages touched by a helper thread while the GPUs
    work (a fresh 8n-byte array costs one page fault per 4 KB on first write; numpy releases the GIL)."""
    import threading

    out = np.empty(int(n), dtype=np.int64)
    th = threading.Thread(target=out.fill, args=(0,), daemon=True)
    th.start()
    return out, th


def distribuited_k_means(data_batch, K, GPU_names, initial_centers=None, n_max_iters=20, *,
                         seed=None, batch_size=None, tol=0.0, empty="keep", path="auto",
                         return_labels=True, out_of_core=None, stream_first_iteration=True):
    """Data-parallel K-Means (Lloyd) on one or more B200s -- drop-in for the reference function.

    Accepts both reference call shapes: ``(data_batch, K, GPU_names, initial_centers,
    n_max_iters)`` (how ``run_experiments`` calls it, :303) and HEAD's ``(data_batch, K,
    GPU_names, n_max_iters)`` (:180), where the centres come from k-means++ (:191; seeded here by
    ``seed``).  ``out_of_core``: keep the rows in host memory and stream them through the GPUs in
    batches of ``batch_size`` rows every iteration (exact Lloyd across batches); None = only when a
    shard does not fit in HBM.  ``GPU_names`` is a list of device names as returned by the reference's
    ``get_available_gpus()`` (or an int = number of GPUs).  ``batch_size`` = rows per pinned H2D
    staging chunk; ``tol`` = optional early stop on the squared centre shift (0 = run all
    ``n_max_iters`` like the reference).  ``stream_first_iteration``: on the tensor path with one GPU per
    process, form the first iteration's sums piece by piece while the rows are still being copied to the GPU.

    Returns the reference's result dictionary (:284-292): ``end_center``, ``cluster_idx``,
    ``centroids`` (labels of the last GPU's shard in the final iteration, before its update), ``init_center``, ``setup_time``,
    ``initialization_time``, ``computation_time``, ``n_iter`` -- plus ``inertia``, ``center_shift``,
    ``converged``.
    """
    torch = _torch()
    setup_ts = time.time()
    if initial_centers is not None and np.isscalar(initial_centers) and not isinstance(initial_centers, bool):
        # HEAD's 4-positional form: the 4th argument is n_max_iters (:180)
        n_max_iters, initial_centers = int(initial_centers), None
    K = int(K)
    n_max_iters = int(n_max_iters)
    if K <= 0 or n_max_iters < 0:
        raise ValueError("K must be positive and n_max_iters non-negative")
    data = _as_host_matrix(data_batch)
    N, D = int(data.shape[0]), int(data.shape[1])
    if K > N:
        raise ValueError(f"K={K} clusters but only N={N} points")
    n_avail = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n_avail == 0:
        raise RuntimeError("no CUDA device: the B200 K-Means path has no CPU fallback")
    dev_ids = parse_gpu_names(GPU_names, n_avail)
    ranges = shard_ranges(N, len(dev_ids))                                    # :184
    on_device = isinstance(data, torch.Tensor) and data.is_cuda

    if initial_centers is None:                                               # :191
        with torch.cuda.device(torch.device("cuda", dev_ids[0])):
            initial_centers, _ = kmeanspp_device(data, K, seed)
    init_center = np.array(initial_centers, dtype=np.float32, copy=True)
    if init_center.shape != (K, D):
        raise ValueError(f"initial_centers shape {init_center.shape} != {(K, D)}")

    if out_of_core is None:
        out_of_core = (not on_device) and not all(
            _fits_in_hbm(r1 - r0, D, K, torch.device("cuda", dev), empty, path) for dev, (r0, r1) in zip(dev_ids, ranges))
    if out_of_core:
        if on_device:
            raise ValueError("out_of_core needs the rows in host memory")
        return _k_means_out_of_core(data, K, dev_ids, ranges, init_center, n_max_iters, batch_size, tol, empty, path,
                                    return_labels, setup_ts)
    shards = [LloydShard(r1 - r0, D, K, torch.device("cuda", dev), empty=empty, path=path)
              for dev, (r0, r1) in zip(dev_ids, ranges)]
    cluster_idx, toucher = _prefaulted_labels(N) if return_labels else (None, None)
    setup_time = float(time.time() - setup_ts)

    # ---- initialisation: H2D of every shard + centroids (:272-274) ----
    initialization_ts = time.time()
    stream_ok = n_max_iters >= 1 and stream_first_iteration and not on_device
    piece = batch_size if batch_size and batch_size >= (1 << 20) else None

    def _load(job):
        s, (r0, r1) = job
        s.set_centroids(init_center)
        if on_device:
            blk = data[r0:r1]
            if blk.device != s.device or blk.dtype != torch.float32 or not blk.is_contiguous():
                blk = blk.to(device=s.device, dtype=torch.float32).contiguous()
            s.bind_points(blk)
            return False
        # the sums of the first iteration are formed piece by piece while the rows are still arriving
        if stream_ok and s.stage_and_accumulate_from_host(data, r0, piece):
            return True
        s.stage_from_host(data, r0, batch_size)
        return False

    if len(shards) == 1:
        flags = [_load((shards[0], ranges[0]))]
    else:
        # one host thread per GPU: the staging calls block on their own copies only (ctypes releases the GIL), so
        # the shards load -- and run their first iteration -- side by side instead of one after the other
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=len(shards)) as pool:
            flags = list(pool.map(_load, zip(shards, ranges)))
    streamed = all(flags)
    if any(flags) and not streamed:
        for s, f in zip(shards, flags):          # mixed outcome (a shard too small to stream): plain first pass there
            if not f:
                s.accumulate()
        streamed = True
    for s in shards:
        torch.cuda.synchronize(s.device)
    initialization_time = float(time.time() - initialization_ts)

    # ---- the Lloyd loop (:277-280) ----
    aux_ts = time.time()
    if streamed and len(shards) == 1:
        n_iter, inertia, shift2 = lloyd_loop(shards, 1, tol, reduce_fn=lambda sh: None, first_accumulated=True)
        if n_max_iters > 1 and not (tol > 0.0 and shift2 <= tol):
            more, inertia, shift2 = shards[0].iterate(n_max_iters - 1, tol)
            n_iter += more
    elif len(shards) == 1:
        n_iter, inertia, shift2 = shards[0].iterate(n_max_iters, tol)
    else:
        n_iter, inertia, shift2 = lloyd_loop(shards, n_max_iters, tol, first_accumulated=streamed)
    for s in shards:
        torch.cuda.synchronize(s.device)
    computation_time = float(time.time() - aux_ts)

    # ---- label pass against the updated centroids (:255, :282) ----
    end_center = shards[0].get_centroids()
    last_labels = None
    if return_labels:
        # labels leave the GPU already widened to the reference's int64, through the shard's pinned staging
        # buffers into one ordinary array (no pinned allocation of N labels inside the call)
        parts = [s.assign() for s in shards]                 # asynchronous on every GPU
        toucher.join()
        if len(shards) == 1:
            shards[0].labels_to_host(parts[0], cluster_idx[ranges[0][0]:ranges[0][1]])
        else:
            # one host thread per GPU again: every shard's labels travel through its own pinned buffers (the C call
            # releases the GIL), so the device-to-host copies of the shards overlap
            from concurrent.futures import ThreadPoolExecutor
            with ThreadPoolExecutor(max_workers=len(shards)) as pool:
                list(pool.map(lambda job: job[0].labels_to_host(job[1], cluster_idx[job[2][0]:job[2][1]]),
                              zip(shards, parts, ranges)))
        if n_iter >= 1:
            # `centroids` (:279, :286): the last GPU's labels of the final iteration, against the centroids
            # BEFORE its update -- not a slice of cluster_idx, which is computed after it (:282)
            last = shards[-1]
            last_labels = last.labels_to_host(last.last_labels(), np.empty(last.n, dtype=np.int64))
    for s in shards:
        s.close()

    return {
        "end_center": end_center,
        "cluster_idx": cluster_idx,
        "centroids": last_labels,
        "init_center": init_center,
        "setup_time": setup_time,
        "initialization_time": initialization_time,
        "computation_time": computation_time,
        "n_iter": n_iter,
        "inertia": inertia,
        "center_shift": float(np.sqrt(shift2)) if shift2 == shift2 else float("nan"),
        "converged": bool(tol > 0.0 and shift2 <= tol),
        # True: iteration 1 ran underneath the H2D copy and is part of initialization_time, not of
        # computation_time (pass stream_first_iteration=False for the reference's timer split, :272-280)
        "first_iteration_streamed": bool(streamed),
    }


class _OutOfCoreShard:
    """One GPU's share of the rows when they do not fit in HBM together: the rows stay in host memory
    and are streamed through one device buffer of `batch_rows` rows.  Presents the interface
    lloyd_loop needs; accumulate() = stage + bind + kmc_accumulate_add for every batch, so several
    batches add up to ONE exact Lloyd iteration (the reference instead clusters every batch on its own
    and averages the end centres, :296-318)."""

    def __init__(self, host, r0, r1, K, device, batch_rows, empty, path):
        torch = _torch()
        self.host, self.r0, self.r1 = host, int(r0), int(r1)
        self.batch_rows = int(min(batch_rows, max(1, r1 - r0)))
        D = int(host.shape[1])
        self.inner = LloydShard(self.batch_rows, D, K, device, empty=empty, path=path)
        with torch.cuda.device(self.inner.device):
            self.buf = torch.empty((self.batch_rows, D), dtype=torch.float32, device=self.inner.device)
        self.device = self.inner.device

    @property
    def partials(self):
        return self.inner.partials

    def _batches(self):
        for b0 in range(self.r0, self.r1, self.batch_rows):
            yield b0, min(self.r1, b0 + self.batch_rows)

    def _load(self, b0, b1):
        self.inner.resize(b1 - b0)
        self.inner.stage_from_host(self.host, b0, None, out=self.buf)

    def set_centroids(self, C):
        self.inner.set_centroids(C)

    def get_centroids(self):
        return self.inner.get_centroids()

    def accumulate(self):
        self.inner.partials.zero_()
        for b0, b1 in self._batches():
            self._load(b0, b1)
            self.inner.accumulate_add()

    def update(self):
        self.inner.update()

    def stats(self):
        return self.inner.stats()

    def assign_to_host(self, out):
        """Labels of rows [r0, r1) into the host int64 array `out` (indexed by global row)."""
        for b0, b1 in self._batches():
            self._load(b0, b1)
            self.inner.labels_to_host(self.inner.assign(), out[b0:b1])

    def close(self):
        self.inner.close()


def _fits_in_hbm(n_rows, D, K, device, empty, path) -> bool:
    """Would a resident shard of n_rows (points + workspace) fit in the free memory of `device`?"""
    torch = _torch()
    lib = L.load()
    ctx = c_void_p()
    flags = L.PATH_FLAGS[path]
    L.check(lib.kmc_create(byref(ctx), torch.device(device).index or 0, int(n_rows), int(D), int(K), flags))
    need = int(lib.kmc_workspace_bytes(ctx)) + int(n_rows) * int(D) * 4
    lib.kmc_destroy(ctx)
    free, _total = torch.cuda.mem_get_info(torch.device(device))
    return need <= 0.92 * free


def distribuited_k_means_spmd(local_rows, K, initial_centers, n_max_iters=20, *, device=None, batch_size=None,
                              tol=0.0, empty="keep", path="auto", return_labels=True, group=None,
                              stream_first_iteration=True):
    """The same Lloyd loop with ONE PROCESS PER GPU (torchrun): every rank passes the rows the
    reference's array_split rule (:184) gives it (`shard_ranges(N, world)[rank]`) and the same
    initial centres; the per-GPU partial sums meet in one NCCL all-reduce per iteration
    (`torch.distributed`, default group or `group`), replacing the reference's add_n on /cpu:0 (:257-258).
    Returns the reference's result keys; `cluster_idx` holds this rank's labels only."""
    torch = _torch()
    import torch.distributed as dist

    setup_ts = time.time()
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device())
    device = torch.device(device)
    rows = _as_host_matrix(local_rows)
    n_local, D = int(rows.shape[0]), int(rows.shape[1])
    init_center = np.array(initial_centers, dtype=np.float32, copy=True)
    shard = LloydShard(n_local, D, int(K), device, empty=empty, path=path)
    labels, toucher = _prefaulted_labels(n_local) if return_labels else (None, None)
    setup_time = float(time.time() - setup_ts)

    initialization_ts = time.time()
    shard.set_centroids(init_center)
    streamed = False
    if isinstance(rows, torch.Tensor) and rows.is_cuda:
        shard.bind_points(rows.to(device=device, dtype=torch.float32).contiguous())
    else:
        if int(n_max_iters) >= 1 and stream_first_iteration:
            streamed = shard.stage_and_accumulate_from_host(rows, 0, batch_size if batch_size and batch_size >= (1 << 20) else None)
        if not streamed:
            shard.stage_from_host(rows, 0, batch_size)
    torch.cuda.synchronize(device)
    initialization_time = float(time.time() - initialization_ts)

    aux_ts = time.time()
    distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    reduce_fn = (lambda s: dist.all_reduce(s[0].partials, group=group)) if distributed else (lambda s: None)
    n_iter, inertia, shift2 = lloyd_loop([shard], int(n_max_iters), tol, reduce_fn=reduce_fn, first_accumulated=streamed)
    torch.cuda.synchronize(device)
    computation_time = float(time.time() - aux_ts)

    end_center = shard.get_centroids()
    last_labels = None
    if return_labels:
        dev_labels = shard.assign()
        toucher.join()
        shard.labels_to_host(dev_labels, labels)
        if n_iter >= 1:
            last_labels = shard.labels_to_host(shard.last_labels(), np.empty(n_local, dtype=np.int64))
    shard.close()
    return {
        "end_center": end_center, "cluster_idx": labels, "centroids": last_labels, "init_center": init_center,
        "setup_time": setup_time, "initialization_time": initialization_time,
        "computation_time": computation_time, "n_iter": n_iter, "inertia": inertia,
        "center_shift": float(np.sqrt(shift2)) if shift2 == shift2 else float("nan"),
        "converged": bool(tol > 0.0 and shift2 <= tol), "first_iteration_streamed": bool(streamed),
    }


def _k_means_out_of_core(data, K, dev_ids, ranges, init_center, n_max_iters, batch_size, tol, empty, path,
                         return_labels, setup_ts):
    torch = _torch()
    N, D = int(data.shape[0]), int(data.shape[1])
    shards = []
    for dev, (r0, r1) in zip(dev_ids, ranges):
        device = torch.device("cuda", dev)
        rows = int(batch_size) if batch_size else (r1 - r0)
        while rows > 1024 and not _fits_in_hbm(rows, D, K, device, empty, path):
            rows = (rows + 1) // 2                      # largest power-of-two fraction of the shard that fits
        shards.append(_OutOfCoreShard(data, r0, r1, K, device, rows, empty, path))
    setup_time = float(time.time() - setup_ts)
    initialization_ts = time.time()
    for s in shards:
        s.set_centroids(init_center)
    initialization_time = float(time.time() - initialization_ts)   # the rows are re-staged every iteration
    aux_ts = time.time()
    n_iter, inertia, shift2 = lloyd_loop(shards, n_max_iters, tol)
    for s in shards:
        torch.cuda.synchronize(s.device)
    computation_time = float(time.time() - aux_ts)
    end_center = shards[0].get_centroids()
    cluster_idx = last_labels = None
    if return_labels:
        cluster_idx = np.empty(N, dtype=np.int64)
        for s in shards:
            s.assign_to_host(cluster_idx)
        # (the pre-update labels of the final iteration are not kept per batch: the slice of cluster_idx stands in)
        last_labels = cluster_idx[ranges[-1][0]:ranges[-1][1]]
    for s in shards:
        s.close()
    return {
        "end_center": end_center, "cluster_idx": cluster_idx, "centroids": last_labels, "init_center": init_center,
        "setup_time": setup_time, "initialization_time": initialization_time, "computation_time": computation_time,
        "n_iter": n_iter, "inertia": inertia,
        "center_shift": float(np.sqrt(shift2)) if shift2 == shift2 else float("nan"),
        "converged": bool(tol > 0.0 and shift2 <= tol), "out_of_core": True,
        "batch_rows": [s.batch_rows for s in shards],
    }


def run_experiments(batches, GPU_names, K, initial_centers, n_max_iters, func=None, **kwargs):
    """The reference's batch driver (:296-318): run ``func`` per batch, average the end centres
    across batches (``np.mean(result, axis=0)``, :310) and add up the three timers."""
    func = func or distribuited_k_means
    result, setup_time, initialization_time, computation_time = [], 0.0, 0.0, 0.0
    for batch in batches:
        batch_run = func(batch, K, GPU_names, initial_centers, n_max_iters, **kwargs)
        result.append(batch_run["end_center"])
        setup_time += batch_run["setup_time"]
        initialization_time += batch_run["initialization_time"]
        computation_time += batch_run["computation_time"]
    return {
        "end_center": np.mean(result, axis=0),
        "init_center": initial_centers,
        "setup_time": setup_time,
        "initialization_time": initialization_time,
        "computation_time": computation_time,
        "n_iter": n_max_iters,
    }
