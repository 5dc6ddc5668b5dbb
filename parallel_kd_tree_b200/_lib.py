"""ctypes binding of libkdtree_b200.so (declarations: include/kdtree_b200.h)."""
from __future__ import annotations

import ctypes as C
import os
import threading

from . import build as _build

_lock = threading.Lock()
_lib = None

c_i32p = C.POINTER(C.c_int32)
c_i64p = C.POINTER(C.c_int64)
c_f32p = C.POINTER(C.c_float)


class KdError(RuntimeError):
    pass


class DistCtx(C.Structure):
    """kd_dist_ctx of include/kdtree_b200.h"""
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("dim", C.c_int), ("with_tree_points", C.c_int),
                ("n_total", C.c_int64), ("heap", C.c_void_p * 8), ("mc_heap", C.c_void_p)]


def _declare(lib):
    vp, i64, i32, f32 = C.c_void_p, C.c_int64, C.c_int, C.c_float
    sigs = {
        "kd_last_error": (C.c_char_p, []),
        "kd_version": (i32, []),
        "kd_free": (None, [vp]),
        "kd_generate_uniform": (i32, [i32, i32, i64, vp]),
        "kd_generate_uniform_rows": (i32, [i32, i32, i64, i64, vp]),
        "kd_generator_selftest": (i64, [i32, i64]),
        "kd_generate_clustered": (i32, [i32, i32, i64, i32, f32, vp]),
        "kd_build_cpu": (i32, [vp, i64, i32, vp, i32]),
        "kd_build_cpu_sortref": (i32, [vp, i64, i32, vp]),
        "kd_nearest_cpu": (i32, [vp, i64, i32, vp, vp, i64, vp, vp, vp, vp]),
        "kd_materialize_nodes_cpu": (i32, [i64, vp, vp, vp, vp]),
        "kd_root_slot": (i64, [i64]),
        "kd_check_tree_cpu": (i32, [vp, i64, i32, vp, vp]),
        "kd_check_tree_ties_cpu": (i32, [vp, i64, i32, vp, vp]),
        "kd_format_float": (i32, [f32, C.c_char_p, i32]),
        "kd_format_result_line": (i32, [i32, f32, C.c_char_p, i32]),
        "kd_format_point": (i32, [vp, i32, i32, C.POINTER(vp), C.POINTER(C.c_size_t)]),
        "kd_format_tree": (i32, [vp, i64, i32, vp, i32, C.POINTER(vp), C.POINTER(C.c_size_t)]),
        "kd_format_head_and_leaves": (i32, [vp, i64, i32, vp, i32, C.POINTER(vp), C.POINTER(C.c_size_t)]),
        "kd_build_workspace_bytes": (C.c_size_t, [i64, i32]),
        "kd_build_gpu": (i32, [vp, i64, i32, vp, vp, vp, C.c_size_t, vp]),
        "kd_debug_enabled": (i32, []),
        "kd_debug_bounds_status": (i32, []),
        "kd_build_gpu_noalloc": (i32, [vp, i64, i32, vp, vp, vp]),
        "kd_build_status_noalloc": (i32, [i64, i32, vp]),
        "kd_build_status": (i32, [vp, i64, i32, vp]),
        "kd_last_launch_count": (i64, []),
        "kd_nearest_gpu": (i32, [vp, vp, i64, i32, vp, i64, vp, vp, vp, vp, vp]),
        "kd_nearest_dense_workspace_bytes": (C.c_size_t, [i64, i64]),
        "kd_nearest_dense_gpu": (i32, [vp, vp, i64, i32, vp, i64, vp, vp, vp, vp, vp, C.c_size_t, vp]),
        "kd_query_order_gpu": (i32, [vp, i64, i32, vp, i64, i32, vp, vp, vp, vp]),
        "kd_nearest_ordered_gpu": (i32, [vp, vp, i64, i32, vp, i64, vp, vp, vp, vp, vp, vp]),
        "kd_knn_ordered_gpu": (i32, [vp, vp, i64, i32, vp, i64, i32, vp, vp, vp, vp]),
        "kd_knn_cpu": (i32, [vp, i64, i32, vp, vp, i64, i32, vp, vp]),
        "kd_knn_gpu": (i32, [vp, vp, i64, i32, vp, i64, i32, vp, vp, vp]),
        "kd_gather_rows_gpu": (i32, [vp, vp, i64, i32, vp, vp]),
        "kd_materialize_nodes_gpu": (i32, [i64, vp, vp, vp, vp, vp]),
        "kd_bbox_gpu": (i32, [vp, i64, i32, vp, vp, vp]),
        "kd_dist_heap_bytes": (C.c_size_t, [i64, i32, i32, i32]),
        "kd_dist_offsets": (i32, [i64, i32, i32, i32, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
        "kd_dist_init": (i32, [C.POINTER(DistCtx), vp, i64, i64, vp]),
        "kd_dist_run": (i32, [C.POINTER(DistCtx), C.c_uint64, vp]),
        "kd_dist_phase": (i32, [C.POINTER(DistCtx), i32, i32, i32, C.c_uint64, vp]),
        "kd_dist_level_counts": (i32, [C.POINTER(DistCtx), C.POINTER(C.c_uint32), vp]),
        "kd_dist_finish": (i32, [C.POINTER(DistCtx), vp]),
        "kd_dist_schedule": (i32, [C.POINTER(C.c_int), C.POINTER(C.c_int), i32]),
        "kd_dist_local_build": (i32, [C.POINTER(DistCtx), vp]),
        "kd_dist_done": (i32, [C.POINTER(DistCtx), i32, C.c_uint64, vp]),
        "kd_dist_status": (i32, [C.POINTER(DistCtx), vp]),
        "kd_forest_buf_bytes": (C.c_size_t, [i32]),
        "kd_forest_min_reduce": (i32, [vp, vp, i32, i32, vp, i32, vp, C.c_uint64, vp, vp]),
        "kd_peer_minmax_buf_bytes": (C.c_size_t, [i32]),
        "kd_peer_minmax_reduce": (i32, [C.POINTER(vp), i32, i32, vp, vp, i32, vp, vp, C.c_uint64, vp, vp]),
        "kd_symm_alloc": (i32, [C.c_size_t, C.POINTER(vp), vp]),
        "kd_symm_open": (i32, [vp, C.POINTER(vp)]),
        "kd_symm_close": (i32, [vp]),
        "kd_symm_free": (i32, [vp]),
        "kd_symm_memset": (i32, [vp, C.c_size_t, vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)   # a missing symbol is a build error, fail loudly
        fn.restype = res
        fn.argtypes = args
    lib._kd_declared = set(sigs)


def lib():
    """Load (building it first if the sources changed) the native library."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                debug = os.environ.get("KD_B200_DEBUG") == "1"     # the -DKD_DEBUG variant (bounds assertions)
                path = _build.LIB_DBG if debug else _build.LIB
                if os.environ.get("KD_B200_NO_BUILD") != "1":
                    path = _build.build_library(debug=debug)
                elif not path.exists():
                    raise KdError(f"{path} is missing and KD_B200_NO_BUILD=1 forbids building it")
                handle = C.CDLL(str(path))
                _declare(handle)
                _lib = handle
    return _lib


def check(status: int, what: str = "") -> None:
    if status != 0:
        msg = lib().kd_last_error()
        raise KdError(f"{what}: {msg.decode() if msg else 'error ' + str(status)}")


def take_string(ptr: C.c_void_p, length: C.c_size_t) -> str:
    try:
        return C.string_at(ptr.value, length.value).decode()
    finally:
        lib().kd_free(ptr)
