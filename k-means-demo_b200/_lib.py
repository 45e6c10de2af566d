"""ctypes binding of libkmeans.so — one line per prototype in include/kmeans.h.

The product path has no fallback: if the CUDA library cannot be loaded, importing the
compute entry points raises; nothing here ever routes to the CPU oracle.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libkmeans.so")

OK = 0
ERR_INVALID_ARG, ERR_CUDA, ERR_NCCL, ERR_OOM, ERR_STATE, ERR_NONFINITE, ERR_UNSUPPORTED = -1, -2, -3, -4, -5, -6, -7
DTYPE_F32, DTYPE_F64 = 0, 1
PATH_AUTO, PATH_DIRECT_F64, PATH_FILTER_F32, PATH_TENSOR = 0, 1, 2, 3
UPDATE_INCREMENTAL, UPDATE_FULL = 0, 1
ABI_VERSION = 4


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("device", C.c_int32), ("path", C.c_int32),
                ("update_mode", C.c_int32), ("exact_inertia", C.c_int32), ("profile_kernels", C.c_int32),
                ("stream", C.c_void_p), ("no_prune", C.c_int32), ("prune_min_points", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("converged", C.c_int32), ("changed", C.c_int64),
                ("inertia", C.c_double), ("rescored", C.c_int64), ("n_total", C.c_int64),
                ("device_ms", C.c_float), ("path_used", C.c_int32)]

    def as_dict(self):
        return {f: getattr(self, f) for f, _ in self._fields_}


class NoAllocResult(C.Structure):
    _fields_ = [("changed", C.c_int64), ("inertia", C.c_double), ("n", C.c_int64), ("status", C.c_int32),
                ("pad", C.c_int32)]


# name -> (restype, argtypes); tests check this table against include/kmeans.h
_vp, _i32, _i64, _u64, _dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_double
PROTOTYPES = {
    "km_abi_version": (_i32, []),
    "km_create": (_i32, [C.POINTER(Config), C.POINTER(_vp)]),
    "km_destroy": (_i32, [_vp]),
    "km_last_error": (C.c_char_p, [_vp]),
    "km_host_alloc": (_i32, [C.POINTER(_vp), _u64]),
    "km_host_free": (_i32, [_vp]),
    "km_comm_unique_id": (_i32, [_vp]),
    "km_comm_init": (_i32, [_vp, _vp, _i32, _i32]),
    "km_upload_points": (_i32, [_vp, _vp, _i64, _i32, _i32]),
    "km_attach_points": (_i32, [_vp, _vp, _i64, _i32, _i32]),
    "km_generate_blobs": (_i32, [_vp, _i64, _i64, _i32, _i32, _u64, _dbl]),
    "km_get_points": (_i32, [_vp, _vp, _i64, _i64]),
    "km_set_centroids": (_i32, [_vp, _vp, _i32]),
    "km_init_forgy": (_i32, [_vp, _i32, _u64]),
    "km_init_kmeanspp": (_i32, [_vp, _i32, _u64]),
    "km_assign": (_i32, [_vp, C.POINTER(Stats)]),
    "km_update": (_i32, [_vp, C.POINTER(Stats)]),
    "km_step": (_i32, [_vp, C.POINTER(Stats)]),
    "km_step_async": (_i32, [_vp]),
    "km_stats_fetch": (_i32, [_vp, C.POINTER(Stats)]),
    "km_run": (_i32, [_vp, _i32, C.POINTER(Stats)]),
    "km_reset_labels": (_i32, [_vp]),
    "km_set_update_mode": (_i32, [_vp, _i32]),
    "km_get_labels_u8": (_i32, [_vp, _vp, _i64]),
    "km_comm_info": (_i32, [_vp, C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_i32)]),
    "km_get_shape": (_i32, [_vp, C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i32), C.POINTER(_i32)]),
    "km_get_labels": (_i32, [_vp, _vp, _i64]),
    "km_get_centroids": (_i32, [_vp, _vp]),
    "km_get_counts": (_i32, [_vp, _vp]),
    "km_get_balance": (_i32, [_vp, C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_dbl)]),
    "km_get_cluster_inertia": (_i32, [_vp, _vp]),
    "km_get_sums": (_i32, [_vp, _vp, _vp, _vp]),
    "km_lloyd_step_noalloc": (_i32, [_i32, _vp, _vp, _i64, _i32, _vp, _i32, _vp, _vp, _vp, _vp]),
    "km_sync": (_i32, [_vp]),
    "km_kernel_time": (_i32, [_vp, C.POINTER(_dbl), C.POINTER(_i32), _i32]),
    "km_set_profile": (_i32, [_vp, _i32]),
    "km_launch_count": (_i64, [_vp]),
}

_lib = None


def load():
    """dlopen libkmeans.so (building it first if the tree has sources but no binary)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        from . import build as _build
        _build.build()
    L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(L, name)  # AttributeError here = the .so does not export what kmeans.h declares
        fn.restype = res
        fn.argtypes = args
    if L.km_abi_version() != ABI_VERSION:
        raise ImportError(f"libkmeans.so ABI {L.km_abi_version()} != binding ABI {ABI_VERSION}")
    _lib = L
    return L
