"""ctypes binding of include/uvaia_b200.h (the C ABI of libuvaia_b200.so).  No torch types cross this boundary.

The library is built in-tree by `make -C uvaia_b200/csrc` (or __graft_entry__.build()).  There is no fallback: if the
shared object is missing this module raises at import, and uvg_create() fails without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("UVG_LIB") or os.path.join(_HERE, "lib", "libuvaia_b200.so")     # UVG_LIB: kernel-variant experiments

UVG_ABI_VERSION = 3


class uvg_item(C.Structure):
    _fields_ = [("score", C.c_int32 * 6), ("ref", C.c_int64)]


class uvg_params(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("nchar", C.c_int32), ("trim", C.c_int32), ("nbest", C.c_int32),
                ("acgt_mode", C.c_int32), ("pool", C.c_int32), ("device", C.c_int32), ("first_round", C.c_int32),
                ("max_round", C.c_int32), ("early_abort", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32),
                ("pair_kernel", C.c_int32), ("closed_form", C.c_int32), ("db_capacity", C.c_int64),
                ("nominal_tail", C.c_int32), ("reserved2", C.c_int32)]


class uvg_ball_params(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("nchar", C.c_int32), ("trim", C.c_int32), ("dist", C.c_int32),
                ("acgt_mode", C.c_int32), ("device", C.c_int32)]


class uvg_query_info(C.Structure):
    _fields_ = [("n_query", C.c_int32), ("nchar", C.c_int32), ("trim", C.c_int32),
                ("n_idx", C.c_int32), ("n_idx_m", C.c_int32), ("n_idx_c", C.c_int32)]


class uvg_stats(C.Structure):
    _fields_ = [("n_refs", C.c_int64), ("n_rounds", C.c_int64), ("n_admitted", C.c_int64),
                ("n_h1_flags", C.c_int64), ("pair_kernel_launches", C.c_int64),
                ("other_kernel_launches", C.c_int64), ("pair_kernel_ms", C.c_double), ("replay_ms", C.c_double),
                ("search_ms", C.c_double), ("encode_ms", C.c_double), ("warp_chunks_executed", C.c_int64),
                ("max_mismatch", C.c_int64), ("pivot_fix_entries", C.c_int64),
                ("set_queries_host_ms", C.c_double), ("set_queries_device_ms", C.c_double)]


# every symbol include/uvaia_b200.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "uvg_create": (C.c_int, [C.POINTER(uvg_params), C.POINTER(_P)]),
    "uvg_destroy": (None, [_P]),
    "uvg_last_error": (C.c_char_p, [_P]),
    "uvg_abi_version": (C.c_int, []),
    "uvg_set_queries": (C.c_int, [_P, _P, C.c_int32, C.POINTER(uvg_query_info)]),
    "uvg_prune_queries": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, C.POINTER(C.c_int32)]),
    "uvg_push_refs": (C.c_int, [_P, _P, C.c_int64, C.c_int32, _P, _P, C.POINTER(C.c_int64)]),
    "uvg_push_refs_device": (C.c_int, [_P, _P, C.c_int64, C.c_int32, _P, _P, C.POINTER(C.c_int64)]),
    "uvg_packed_bytes": (C.c_int, [_P, C.c_int64, C.c_int64, C.POINTER(C.c_size_t)]),
    "uvg_export_packed": (C.c_int, [_P, C.c_int64, C.c_int64, _P, C.c_size_t]),
    "uvg_import_packed": (C.c_int, [_P, _P, C.c_size_t, C.POINTER(C.c_int64)]),
    "uvg_drop_refs": (C.c_int, [_P]),
    "uvg_clear_refs": (C.c_int, [_P]),
    "uvg_db_count": (C.c_int64, [_P]),
    "uvg_mark_file_end": (C.c_int, [_P, C.c_int64]),
    "uvg_search": (C.c_int, [_P]),
    "uvg_search_async": (C.c_int, [_P]),
    "uvg_wait": (C.c_int, [_P]),
    "uvg_reset_search": (C.c_int, [_P]),
    "uvg_get_heap": (C.c_int, [_P, C.c_int32, C.POINTER(uvg_item), C.POINTER(C.c_int32)]),
    "uvg_get_all_heaps": (C.c_int, [_P, _P, _P]),
    "uvg_get_ever_best": (C.c_int, [_P, C.c_int64, C.c_int64, _P]),
    "uvg_get_stats": (C.c_int, [_P, C.POINTER(uvg_stats)]),
    "uvg_shard_of": (C.c_int, [C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int64)]),
    "uvg_nccl_unique_id": (C.c_int, [_P]),
    "uvg_nccl_init": (C.c_int, [_P, _P]),
    "uvg_set_global_count": (C.c_int, [_P, C.c_int64]),
    "uvg_set_search_limit": (C.c_int, [_P, C.c_int64]),
    "uvg_ball_create": (C.c_int, [C.POINTER(uvg_ball_params), C.POINTER(_P)]),
    "uvg_ball_destroy": (None, [_P]),
    "uvg_ball_last_error": (C.c_char_p, [_P]),
    "uvg_ball_radius": (C.c_int32, [_P]),
    "uvg_ball_set_queries": (C.c_int, [_P, _P, C.c_int32, C.POINTER(uvg_query_info)]),
    "uvg_ball_push_refs": (C.c_int, [_P, _P, C.c_int64, C.c_int32, _P, _P]),
    "uvg_ball_classify_device_noalloc": (C.c_int, [_P, _P, C.c_int64, C.c_int32, _P, _P]),
    "uvg_synth_rows_device": (C.c_int, [_P, _P, C.c_int32, C.c_uint64, C.c_int64, C.c_int64, C.c_double, C.c_double, _P]),
    "uvg_device_alloc": (_P, [_P, C.c_size_t]),
    "uvg_device_free": (None, [_P, _P]),
    "uvg_memcpy_d2h": (C.c_int, [_P, _P, _P, C.c_size_t]),
    "uvg_memcpy_h2d": (C.c_int, [_P, _P, _P, C.c_size_t]),
    "uvg_host_alloc_pinned": (_P, [C.c_size_t]),
    "uvg_host_free_pinned": (None, [_P]),
    "uvg_debug_pair_counts": (C.c_int, [_P, C.c_int64, C.c_int64, _P]),
    "uvg_debug_pair_counts_rows": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int64, C.c_int64, _P]),
    "uvg_debug_ref_meta": (C.c_int, [_P, C.c_int64, C.c_int64, _P, _P, _P]),
    "uvg_debug_unpack_ref": (C.c_int, [_P, C.c_int64, _P]),
}


def load(path: str = LIB_PATH) -> C.CDLL:
    if not os.path.exists(path):
        raise OSError(f"{path} is missing: build it with `make -C uvaia_b200/csrc` (there is no CPU fallback)")
    lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.uvg_abi_version() != UVG_ABI_VERSION:
        raise OSError("libuvaia_b200.so ABI version mismatch")
    return lib


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        _lib = load()
    return _lib
