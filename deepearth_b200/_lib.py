"""ctypes binding of libde_grid4d.so (the C ABI declared in include/de_grid4d.h).

There is deliberately no fallback: if the shared library is missing or a call fails, an
exception is raised.  Build it with ``python -m deepearth_b200.build`` (nvcc, sm_100a).
"""
from __future__ import annotations

import ctypes
import os
import threading

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libde_grid4d.so")

DE_OK = 0
ABI_VERSION = 5
DE_F32, DE_F16, DE_F64, DE_I32, DE_BF16 = 0, 1, 2, 3, 4
DE_SORT_NOALLOC_MAX_BUCKETS = 65536
DE_INTERP_NEAREST, DE_INTERP_LINEAR = 0, 1
DE_FLAG_STAGED_BWD, DE_FLAG_SMEM_COARSE, DE_FLAG_NO_L2_HINTS, DE_FLAG_NO_REGROUP = 0x100, 0x200, 0x400, 0x800
DE_FLAG_GRAD_GROUPED = 0x1000
DE_FLAG_NO_SMALL1D = 0x4000
DE_MAX_LEVELS_PER_CALL = 256

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_i64 = ctypes.c_int64
c_u32 = ctypes.c_uint32
c_float = ctypes.c_float


class DeLevelDesc(ctypes.Structure):
    """Mirror of ``de_level_desc`` (include/de_grid4d.h)."""
    _fields_ = [
        ("table", c_void_p),
        ("grad_table", c_void_p),
        ("table_size", c_u32),
        ("resolution", c_u32),
        ("out_col", c_u32),
        ("n_dims", ctypes.c_uint8),
        ("dim", ctypes.c_uint8 * 4),
        ("dense", ctypes.c_uint8),
        ("mask_id", ctypes.c_uint8),
        ("reserved", ctypes.c_uint8),
    ]


class DeFieldDesc(ctypes.Structure):
    """Mirror of ``de_field_desc`` (include/de_grid4d.h)."""
    _fields_ = [("lat_start", ctypes.c_double), ("lat_stop", ctypes.c_double), ("lon_start", ctypes.c_double),
                ("lon_stop", ctypes.c_double), ("altitude", ctypes.c_double), ("time", ctypes.c_double),
                ("n_lat", ctypes.c_int32), ("n_lon", ctypes.c_int32), ("has_time", ctypes.c_int32),
                ("reserved", ctypes.c_int32)]


# name -> (restype, argtypes); every symbol include/de_grid4d.h declares
SIGNATURES = {
    "de_abi_version": (c_int, []),
    "de_last_error": (ctypes.c_char_p, []),
    "de_corners": (c_int, [c_u32, c_int]),
    "de_grid4d_hash_indices": (c_int, [ctypes.POINTER(DeLevelDesc), c_int, c_void_p, c_i64, c_int, c_u32,
                                       c_void_p, c_void_p]),
    "de_grid4d_forward": (c_int, [ctypes.POINTER(DeLevelDesc), c_int, c_void_p, c_i64, c_int, c_int, c_int,
                                  c_float, c_void_p, c_void_p, c_void_p, c_i64, c_u32, c_int, c_void_p]),
    "de_grid4d_backward": (c_int, [ctypes.POINTER(DeLevelDesc), c_int, c_void_p, c_i64, c_int, c_int,
                                   c_void_p, c_void_p, c_void_p, c_i64, c_u32, c_int, c_void_p, ctypes.c_size_t,
                                   c_void_p]),
    "de_grid4d_backward_workspace_bytes": (ctypes.c_size_t, [ctypes.POINTER(DeLevelDesc), c_int, c_int, c_i64, c_u32]),
    "de_grid4d_backward_workspace_bytes_ex": (ctypes.c_size_t, [ctypes.POINTER(DeLevelDesc), c_int, c_int, c_i64, c_u32, c_int]),
    "de_grid4d_backward_group_width": (c_int, [ctypes.POINTER(DeLevelDesc), c_int, c_int, c_i64, c_u32, c_int]),
    "de_grid4d_sort_workspace_bytes": (ctypes.c_size_t, [c_int, c_int]),
    "de_grid4d_sort_points_noalloc": (c_int, [c_void_p, c_i64, c_int, ctypes.POINTER(ctypes.c_uint8), c_int, c_int, c_void_p, c_void_p]),
    "de_grid4d_sort_points": (c_int, [c_void_p, c_i64, c_int, ctypes.POINTER(ctypes.c_uint8), c_int, c_int, c_void_p, c_void_p,
                                      ctypes.c_size_t, c_void_p]),
    "de_grid4d_backward_sorted": (c_int, [ctypes.POINTER(DeLevelDesc), c_int, c_void_p, c_i64, c_int, c_int, c_void_p, c_void_p,
                                          c_void_p, c_i64, c_u32, c_void_p, c_void_p]),
    "de_tables_to_f16": (c_int, [ctypes.POINTER(DeLevelDesc), c_int, c_int, ctypes.POINTER(c_void_p), c_float,
                                 c_void_p]),
    "de_geo2xyzt_norm": (c_int, [c_void_p, c_int, c_void_p, c_int, ctypes.POINTER(ctypes.c_double),
                                 ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double), c_void_p,
                                 c_void_p, c_i64, c_void_p]),
    "de_mlp_head_supported": (c_int, [c_int, c_int, c_int]),
    "de_mlp_head_forward": (c_int, [c_void_p, c_i64, c_int, c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_float,
                                    c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_void_p, c_i64, c_void_p]),
    "de_mlp_head_backward": (c_int, [c_void_p, c_void_p, c_i64, c_int, c_i64, c_i64, c_void_p, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_void_p, c_float, c_int, c_int, c_int, c_int, c_void_p, c_i64, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "de_time_from_components": (c_int, [c_void_p, c_int, c_i64, c_i64, c_int, c_int, c_int, c_int,
                                        ctypes.POINTER(ctypes.c_double), c_void_p, c_void_p, c_void_p]),
    "de_projection_supported": (c_int, [c_int, c_int, c_int]),
    "de_projection_forward": (c_int, [c_void_p, c_i64, c_int, c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_float,
                                      c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_i64, c_void_p]),
    "de_projection_forward_ex": (c_int, [c_void_p, c_i64, c_int, c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_float,
                                         c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_i64, c_void_p]),
    "de_projection_backward_ex": (c_int, [c_void_p, c_void_p, c_int, c_i64, c_int, c_i64, c_i64, c_void_p, c_void_p, c_void_p, c_void_p,
                                          c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_i64, c_void_p, c_void_p,
                                          c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, ctypes.c_size_t, c_void_p]),
    "de_projection_backward_workspace_bytes": (ctypes.c_size_t, [c_i64, c_int, c_int, c_int]),
    "de_projection_backward": (c_int, [c_void_p, c_void_p, c_i64, c_int, c_i64, c_i64, c_void_p, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_i64, c_void_p, c_void_p,
                                       c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, ctypes.c_size_t, c_void_p]),
    "de_gemm_bf16": (c_int, [c_void_p, c_i64, c_void_p, c_i64, c_i64, c_int, c_i64, c_void_p, c_void_p, c_i64, c_int, c_int, c_int,
                             c_void_p]),
    "de_geo_bbox": (c_int, [c_void_p, c_int, c_void_p, c_i64, c_void_p]),
    "de_set_l2_fetch_granularity": (c_int, [c_int]),
    "de_get_l2_fetch_granularity": (c_int, [ctypes.POINTER(c_int)]),
    "de_set_persisting_l2": (c_int, [ctypes.c_longlong]),
    "de_get_persisting_l2": (c_int, [ctypes.POINTER(ctypes.c_longlong), ctypes.POINTER(ctypes.c_longlong)]),
    "de_field_xyzt": (c_int, [ctypes.POINTER(DeFieldDesc), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double),
                              ctypes.POINTER(ctypes.c_double), c_void_p, c_void_p]),
    "de_grid4d_field_forward": (c_int, [ctypes.POINTER(DeLevelDesc), c_int, ctypes.POINTER(DeFieldDesc),
                                        ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double),
                                        ctypes.POINTER(ctypes.c_double), c_int, c_int, c_float, c_void_p, c_i64, c_void_p,
                                        c_u32, c_int, c_void_p]),
    "de_st_embedding_forward": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_int,
                                        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "de_st_embedding_backward": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_i64, c_void_p, c_int, c_void_p, c_void_p,
                                         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                         c_void_p, c_void_p]),
    "de_st_embedding_tc_supported": (c_int, [c_int]),
    "de_st_embedding_tc_workspace_bytes": (ctypes.c_size_t, [c_i64, c_int, c_int]),
    "de_st_embedding_forward_tc": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_int,
                                           c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                           ctypes.c_size_t, c_void_p]),
    "de_st_embedding_backward_tc": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_i64, c_void_p, c_int, c_void_p, c_void_p,
                                            c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                            c_void_p, ctypes.c_size_t, c_void_p]),
    "de_sum_f32": (c_int, [c_void_p, c_i64, c_void_p, c_void_p]),
    "de_probe_gather": (c_int, [c_void_p, c_i64, c_int, c_int, c_i64, c_int, c_int, c_void_p, c_void_p]),
}

_lock = threading.Lock()
_lib = None


class DeepEarthB200Error(RuntimeError):
    pass


def load() -> ctypes.CDLL:
    """Load the shared library once; raise loudly if it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise DeepEarthB200Error(
                f"{LIB_PATH} is missing: the CUDA library has not been built. Run "
                "`python -m deepearth_b200.build` (needs nvcc; targets sm_100a). There is no CPU fallback.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)   # AttributeError here = header / library mismatch
            fn.restype = res
            fn.argtypes = args
        if lib.de_abi_version() != ABI_VERSION:
            raise DeepEarthB200Error(f"ABI version {lib.de_abi_version()} != {ABI_VERSION}: rebuild libde_grid4d.so")
        _lib = lib
    return _lib


def check(rc: int, what: str) -> None:
    if rc != DE_OK:
        msg = load().de_last_error().decode("utf-8", "replace")
        raise DeepEarthB200Error(f"{what} failed (code {rc}): {msg}")
