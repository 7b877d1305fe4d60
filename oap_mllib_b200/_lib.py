"""ctypes binding of libkmeans_b200.so (the C ABI in include/kmb200.h).  Fails loudly: a missing
library raises at first use -- nothing in this package computes on the CPU instead."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))


def library_path() -> str:
    return os.environ.get("KMB200_LIBRARY", os.path.join(HERE, "libkmeans_b200.so"))


class Kmb200Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"kmb200 error {code}: {msg}")
        self.code = code


class Profile(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("kernel_launches", C.c_int32), ("rechecked_rows", C.c_int64),
                ("empty_cluster_events", C.c_int32), ("full_rechecks", C.c_int64), ("total_ms", C.c_float), ("assign_ms", C.c_float),
                ("recheck_ms", C.c_float), ("bucket_ms", C.c_float), ("accumulate_ms", C.c_float),
                ("allreduce_ms", C.c_float), ("finalize_ms", C.c_float), ("assign_main_ms", C.c_float), ("coarse_undecided_rows", C.c_int64),
                ("cascade_dropped_at", C.c_int32), ("two_level_dropped_at", C.c_int32), ("assign_variant", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


# every symbol include/kmb200.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SIGNATURES = {
    "kmb200_device_ok": (C.c_int, []),
    "kmb200_device_count": (C.c_int, []),
    "kmb200_set_device": (C.c_int, [C.c_int32]),
    "kmb200_table_create": (C.c_int, [C.c_int64, C.c_int32, C.POINTER(_P)]),
    "kmb200_table_append_f64": (C.c_int, [_P, C.c_int64, _P, C.c_int64, C.c_int32]),
    "kmb200_table_append_f32": (C.c_int, [_P, C.c_int64, _P, C.c_int64, C.c_int32]),
    "kmb200_table_append_columns_f64": (C.c_int, [_P, C.c_int64, _P, C.c_int64, C.c_int32]),
    "kmb200_table_append_columns_f32": (C.c_int, [_P, C.c_int64, _P, C.c_int64, C.c_int32]),
    "kmb200_table_append_csr_f64": (C.c_int, [_P, C.c_int64, C.c_int64, _P, _P, _P]),
    "kmb200_table_truncate": (C.c_int, [_P, C.c_int64]),
    "kmb200_pinned_alloc": (C.c_int, [C.c_size_t, C.POINTER(_P)]),
    "kmb200_pinned_free": (C.c_int, [_P]),
    "kmb200_table_read_rows": (C.c_int, [_P, C.c_int64, C.c_int64, _P]),
    "kmb200_table_rows": (C.c_int64, [_P]),
    "kmb200_table_cols": (C.c_int32, [_P]),
    "kmb200_table_fill_blobs": (C.c_int, [_P, _P, C.c_int32, C.c_uint64, C.c_int64]),
    "kmb200_table_free": (C.c_int, [_P]),
    "kmb200_comm_unique_id": (C.c_int, [_P]),
    "kmb200_comm_init": (C.c_int, [C.c_int32, C.c_int32, _P, C.POINTER(_P)]),
    "kmb200_comm_rank": (C.c_int32, [_P]),
    "kmb200_comm_size": (C.c_int32, [_P]),
    "kmb200_comm_destroy": (C.c_int, [_P]),
    "kmb200_fit": (C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, C.c_double, C.c_int32, _P,
                             C.POINTER(C.c_double), C.POINTER(C.c_int32)]),
    "kmb200_assign": (C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, _P, _P, C.POINTER(C.c_double)]),
    "kmb200_get_profile": (C.c_int, [C.POINTER(Profile)]),
    "kmb200_set_option": (C.c_int, [C.c_char_p, C.c_double]),
    "kmb200_cluster_sizes_noalloc": (C.c_int, [_P, C.c_int64, C.c_int32, _P, _P]),
    "kmb200_debug_scores": (C.c_int, [_P, _P, C.c_int32, C.c_int32, _P, C.POINTER(C.c_int32)]),
    "kmb200_init_costs": (C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_double)]),
    "kmb200_init_sample": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, C.c_double, _P, C.c_int32, C.c_int32, _P, C.c_int64,
                                    C.POINTER(C.c_int64)]),
    "kmb200_xorshift_next_int": (C.c_int32, [C.c_int64]),
    "kmb200_xorshift_next_double": (C.c_double, [C.c_int64, C.c_int32]),
    "kmb200_pca_train": (C.c_int, [_P, _P, C.c_int32, _P, _P]),
    "kmb200_pca_gram": (C.c_int, [_P, _P, C.c_int32, _P, _P, _P]),
    "kmb200_sym_eig": (C.c_int, [_P, C.c_int32, _P, _P]),
    "kmb200_debug_dump": (C.c_int, [_P, C.c_char_p, C.c_size_t, _P, C.c_size_t]),
    "kmb200_last_error": (C.c_char_p, []),
    "kmb200_version": (C.c_char_p, []),
}

_lib = None


def lib() -> C.CDLL:
    """The loaded CUDA library.  Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        path = library_path()
        if not os.path.exists(path):
            raise Kmb200Error(-2, f"{path} not found: build it with `python -m oap_mllib_b200.build` "
                                  "(nvcc, sm_100a). There is no CPU fallback.")
        handle = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the library does not export it
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise Kmb200Error(rc, lib().kmb200_last_error().decode("utf-8", "replace"))


def profile() -> dict:
    p = Profile()
    check(lib().kmb200_get_profile(C.byref(p)))
    return p.as_dict()
