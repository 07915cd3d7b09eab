"""ctypes binding of libkmc_b200.so (the C ABI declared in include/kmc.h).

The library is the product: if it is missing or a symbol cannot be resolved this module raises --
there is no Python / PyTorch / CPU fallback behind it.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_uint32, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libkmc_b200.so")

# constants mirrored from include/kmc.h
KMC_EMPTY_KEEP, KMC_EMPTY_NAN = 0x0, 0x1
KMC_PATH_AUTO, KMC_PATH_SMALL, KMC_PATH_SIMT, KMC_PATH_TC = 0x00, 0x10, 0x20, 0x30
KMC_DTYPE_F32, KMC_DTYPE_F64 = 0, 1
KMC_STAT_INERTIA, KMC_STAT_SHIFT2, KMC_STAT_AMBIGUOUS, KMC_STAT_KERNEL_PATH, KMC_STAT_LAUNCHES = 0, 1, 2, 3, 4
KMC_STAT_X_SCALED_MAX = 5
KMC_STAT_CHANGED, KMC_STAT_FULL_PASS = 6, 7
KMC_STAT_WALK_OVERFLOW = 8
KMC_STAT_COUNT = 12
KMC_PARAM_TC_GAMMA = 1
KMC_PARAM_PROFILE = 2
KMC_PARAM_TC_TERMS = 3
KMC_PARAM_REFINE_KERNEL = 4
KMC_PARAM_CHUNKS = 5
KMC_PARAM_INCREMENTAL = 6
KMC_PARAM_TC_PAIR = 7
KMC_PARAM_TC_TILE_N = 8
KMC_PARAM_DELTA_ATOMIC_ROWS = 9
KMC_PARAM_TC_ACC16 = 10
KMC_PHASE_COUNT = 5
PATH_NAMES = {KMC_PATH_SMALL: "small", KMC_PATH_SIMT: "simt", KMC_PATH_TC: "tc"}
PATH_FLAGS = {"auto": KMC_PATH_AUTO, "small": KMC_PATH_SMALL, "simt": KMC_PATH_SIMT, "tc": KMC_PATH_TC}

# every symbol include/kmc.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "kmc_version": (c_int, []),
    "kmc_last_error": (c_char_p, []),
    "kmc_create": (c_int, [POINTER(c_void_p), c_int, c_int64, c_int64, c_int64, c_uint32]),
    "kmc_destroy": (None, [c_void_p]),
    "kmc_workspace_bytes": (c_int64, [c_void_p]),
    "kmc_bind_workspace": (c_int, [c_void_p, c_void_p, c_int64]),
    "kmc_bind_points": (c_int, [c_void_p, c_void_p, c_void_p]),
    "kmc_stage_points": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int64, c_void_p]),
    "kmc_set_centroids": (c_int, [c_void_p, c_void_p, c_void_p]),
    "kmc_get_centroids": (c_int, [c_void_p, c_void_p, c_void_p]),
    "kmc_centroids_ptr": (c_void_p, [c_void_p]),
    "kmc_accumulate": (c_int, [c_void_p, c_void_p]),
    "kmc_partials_ptr": (c_void_p, [c_void_p]),
    "kmc_partials_len": (c_int64, [c_void_p]),
    "kmc_resize": (c_int, [c_void_p, c_int64]),
    "kmc_accumulate_add": (c_int, [c_void_p, c_void_p]),
    "kmc_update": (c_int, [c_void_p, c_void_p]),
    "kmc_iterate": (c_int, [c_void_p, c_int, c_double, POINTER(c_int), POINTER(c_double), POINTER(c_double), c_void_p]),
    "kmc_bind_points_streamed": (c_int, [c_void_p, c_void_p, c_void_p]),
    "kmc_accumulate_rows": (c_int, [c_void_p, c_int64, c_int64, c_void_p]),
    "kmc_assign": (c_int, [c_void_p, c_void_p, c_void_p]),
    "kmc_last_labels": (c_int, [c_void_p, c_void_p, c_void_p]),
    "kmc_gather_centroids": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "kmc_labels_to_host": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p]),
    "kmc_read_stats": (c_int, [c_void_p, POINTER(c_double), c_void_p]),
    "kmc_set_param": (c_int, [c_void_p, c_int, c_double]),
    "kmc_get_phase_ms": (c_int, [c_void_p, POINTER(c_float), c_int]),
    "kmc_debug_tc_scores": (c_int, [c_void_p, c_void_p, c_void_p]),
    "kmc_kmeanspp_scratch_bytes": (c_int64, [c_int64, c_int64]),
    "kmc_kmeanspp": (c_int, [c_int, c_void_p, c_int64, c_int64, c_int64, ctypes.c_uint64, c_void_p, c_void_p, c_void_p,
                             c_int64, c_void_p]),
    "kmc_cluster_sums_noalloc": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64,
                                         c_void_p, c_void_p]),
}

_lib = None


class KmcError(RuntimeError):
    """A CUDA (code > 0) or library (code < 0) failure reported by libkmc_b200."""

    def __init__(self, code: int, message: str):
        super().__init__(f"libkmc_b200 error {code}: {message}")
        self.code = code


def load() -> ctypes.CDLL:
    """Load the shared library and bind every declared symbol (raises if anything is missing)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no fallback implementation)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so does not export it
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(code: int) -> None:
    if code != 0:
        msg = load().kmc_last_error()
        raise KmcError(code, msg.decode("utf-8", "replace") if msg else "")
