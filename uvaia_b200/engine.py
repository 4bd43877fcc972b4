"""Thin object wrapper over the C ABI (uvaia_b200/_capi.py): numpy arrays in, numpy arrays out.

Mirrors the seam of the reference that the CUDA layer replaces (src/nearest.c): `set_queries` ~ create_query_indices,
`push_refs` ~ the batch fill with quick_count_sequence_non_N + ref filter (nearest.c:250-286), `search` ~ the two
OpenMP loops (nearest.c:293-299), `heap` ~ heap_finalise_heap_qsort + save_distance_table reads.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _capi


PAIR_KERNELS = {"auto": 0, "general": 1, "pivot": 2}


class UvgError(RuntimeError):
    pass


def prune_queries(text: np.ndarray, trim=0, acgt=False, keep_more_resolved=True):
    """-k / --keep_resolved (exclude_redundant_query_sequences, src/fastaseq.c:797-841): boolean mask of the ordered
    query rows that stay; the caller passes text[mask] to Engine.set_queries (nearest.c:219-224)."""
    text = np.ascontiguousarray(text, dtype=np.uint8)
    assert text.ndim == 2
    keep = np.empty(text.shape[0], np.uint8)
    n = C.c_int32(0)
    rc = _capi.lib().uvg_prune_queries(text.ctypes.data, text.shape[0], text.shape[1], int(trim), int(bool(acgt)),
                                       int(bool(keep_more_resolved)), keep.ctypes.data, C.byref(n))
    if rc != 0:
        raise UvgError(f"uvg_prune_queries failed ({rc})")
    assert int(keep.sum()) == n.value
    return keep.astype(bool)


class Ball:
    """uvaiaball on the GPU (uvg_ball_*, csrc/ball.cu): which reference rows are within `dist` differences of some query"""

    def __init__(self, nchar, dist=5, trim=0, acgt=False, device=0):
        self._lib = _capi.lib()
        p = _capi.uvg_ball_params(_capi.UVG_ABI_VERSION, int(nchar), int(trim), int(dist), int(bool(acgt)), int(device))
        h = C.c_void_p()
        rc = self._lib.uvg_ball_create(C.byref(p), C.byref(h))
        if rc != 0:
            raise UvgError(f"uvg_ball_create failed ({rc}): {self._lib.uvg_ball_last_error(None).decode()}")
        self._h = h
        self.nchar = int(nchar)
        self.dist = int(self._lib.uvg_ball_radius(h))

    def _ck(self, rc, what):
        if rc != 0:
            raise UvgError(f"{what} failed ({rc}): {self._lib.uvg_ball_last_error(self._h).decode()}")

    def close(self):
        if getattr(self, "_h", None):
            self._lib.uvg_ball_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_queries(self, text: np.ndarray):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        assert text.ndim == 2 and text.shape[1] == self.nchar
        info = _capi.uvg_query_info()
        self._ck(self._lib.uvg_ball_set_queries(self._h, text.ctypes.data, text.shape[0], C.byref(info)), "uvg_ball_set_queries")
        return info

    def push_refs(self, text: np.ndarray, min_non_n=0):
        """-> (keep uint8[n]: 0 too ambiguous / 1 outside / 2 inside the ball, non_n int32[n])"""
        text = np.ascontiguousarray(text, dtype=np.uint8)
        n = text.shape[0]
        keep, nn = np.zeros(n, np.uint8), np.zeros(n, np.int32)
        if n:
            assert text.ndim == 2 and text.shape[1] == self.nchar
            self._ck(self._lib.uvg_ball_push_refs(self._h, text.ctypes.data, n, int(min_non_n), keep.ctypes.data, nn.ctypes.data), "uvg_ball_push_refs")
        return keep, nn


    NOALLOC_MAX_ROWS = 65536                                  # UVG_BALL_NOALLOC_MAX_ROWS (include/uvaia_b200.h)

    def classify_device_noalloc(self, d_text, n, min_non_n, d_keep, d_non_n=0):
        """uvg_ball_classify_device_noalloc: rows and results are DEVICE addresses of the caller (ints); the call allocates
        nothing (static device scratch, at most NOALLOC_MAX_ROWS rows per call)"""
        self._ck(self._lib.uvg_ball_classify_device_noalloc(self._h, int(d_text), int(n), int(min_non_n), int(d_keep), int(d_non_n) or None),
                 "uvg_ball_classify_device_noalloc")


class Engine:
    def __init__(self, nchar, nbest=100, trim=0, acgt=False, pool=512, device=0, first_round=0, max_round=0,
                 early_abort=False, rank=0, world=1, db_capacity=0, pair_kernel=None, closed_form=False, nominal_tail=False):
        if pair_kernel is None:      # test / debug knob: UVG_PAIR_KERNEL=auto|general|pivot (include/uvaia_b200.h, uvg_params.pair_kernel)
            pair_kernel = PAIR_KERNELS[os.environ.get("UVG_PAIR_KERNEL", "auto")]
        self._lib = _capi.lib()
        p = _capi.uvg_params(_capi.UVG_ABI_VERSION, int(nchar), int(trim), int(nbest), int(bool(acgt)), int(pool),
                             int(device), int(first_round), int(max_round), int(bool(early_abort)), int(rank), int(world),
                             int(pair_kernel), int(bool(closed_form)), int(db_capacity), int(bool(nominal_tail)), 0)
        h = C.c_void_p()
        rc = self._lib.uvg_create(C.byref(p), C.byref(h))
        if rc != 0:
            raise UvgError(f"uvg_create failed ({rc}): {self._lib.uvg_last_error(None).decode()}")
        self._h = h
        self.nchar = int(nchar)
        self.K = max(2, int(nbest))
        self.acgt = bool(acgt)
        self.n_query = 0
        self.info = None
        self.rank, self.world = int(rank), int(world)
        self.first_round, self.max_round = int(first_round), int(max_round)

    # ---- plumbing ----
    def _ck(self, rc, what):
        if rc != 0:
            raise UvgError(f"{what} failed ({rc}): {self._lib.uvg_last_error(self._h).decode()}")

    def close(self):
        if getattr(self, "_h", None):
            self._lib.uvg_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- queries ----
    def set_queries(self, text: np.ndarray):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        assert text.ndim == 2 and text.shape[1] == self.nchar
        info = _capi.uvg_query_info()
        self._ck(self._lib.uvg_set_queries(self._h, text.ctypes.data, text.shape[0], C.byref(info)), "uvg_set_queries")
        self.n_query = text.shape[0]
        self.info = info
        return info

    # ---- references ----
    def push_refs(self, text: np.ndarray, min_non_n=0):
        """-> (non_n int32[n], accepted uint8[n], n_accepted)"""
        text = np.ascontiguousarray(text, dtype=np.uint8)
        if text.size == 0:
            return np.zeros(0, np.int32), np.zeros(0, np.uint8), 0
        assert text.ndim == 2 and text.shape[1] == self.nchar
        n = text.shape[0]
        nn = np.empty(n, np.int32)
        acc = np.empty(n, np.uint8)
        na = C.c_int64(0)
        self._ck(self._lib.uvg_push_refs(self._h, text.ctypes.data, n, int(min_non_n), nn.ctypes.data, acc.ctypes.data,
                                         C.byref(na)), "uvg_push_refs")
        return nn, acc, na.value

    def push_refs_ptr(self, host_ptr, n, min_non_n=0):
        """host buffer given as a raw address (e.g. pinned memory); returns n_accepted"""
        na = C.c_int64(0)
        self._ck(self._lib.uvg_push_refs(self._h, host_ptr, int(n), int(min_non_n), None, None, C.byref(na)), "uvg_push_refs")
        return na.value

    def push_refs_device(self, d_ptr, n, min_non_n=0, want_outputs=False):
        na = C.c_int64(0)
        if want_outputs:
            nn = np.empty(n, np.int32)
            acc = np.empty(n, np.uint8)
            self._ck(self._lib.uvg_push_refs_device(self._h, d_ptr, int(n), int(min_non_n), nn.ctypes.data, acc.ctypes.data,
                                                    C.byref(na)), "uvg_push_refs_device")
            return nn, acc, na.value
        self._ck(self._lib.uvg_push_refs_device(self._h, d_ptr, int(n), int(min_non_n), None, None, C.byref(na)),
                 "uvg_push_refs_device")
        return na.value

    # ---- packed reference cache ----
    def export_packed(self, first=0, n=None, out=None) -> np.ndarray:
        """the resident packed references [first, first+n) as one relocatable blob (uint8 array); `out` may be a
        preallocated (e.g. pinned) uint8 array"""
        n = self.db_count - first if n is None else n
        nb = C.c_size_t(0)
        self._ck(self._lib.uvg_packed_bytes(self._h, int(first), int(n), C.byref(nb)), "uvg_packed_bytes")
        blob = np.empty(nb.value, np.uint8) if out is None else out[:nb.value]
        assert blob.size == nb.value and blob.flags.c_contiguous
        self._ck(self._lib.uvg_export_packed(self._h, int(first), int(n), blob.ctypes.data, blob.size), "uvg_export_packed")
        return blob

    def import_packed(self, blob) -> int:
        """append a blob made by export_packed (uint8 array or a raw (address, bytes) pair); -> references appended"""
        n = C.c_int64(0)
        if isinstance(blob, tuple):
            addr, size = blob
        else:
            blob = np.ascontiguousarray(blob, dtype=np.uint8)
            addr, size = blob.ctypes.data, blob.size
        self._ck(self._lib.uvg_import_packed(self._h, addr, size, C.byref(n)), "uvg_import_packed")
        return n.value

    def mark_file_end(self, n_accepted_so_far=-1):
        """the references accepted so far end a -r file (the reference's batch ends with its file, nearest.c:245-286)"""
        self._ck(self._lib.uvg_mark_file_end(self._h, int(n_accepted_so_far)), "uvg_mark_file_end")

    def drop_refs(self):
        self._ck(self._lib.uvg_drop_refs(self._h), "uvg_drop_refs")

    def clear_refs(self):
        self._ck(self._lib.uvg_clear_refs(self._h), "uvg_clear_refs")

    @property
    def db_count(self):
        return int(self._lib.uvg_db_count(self._h))

    # ---- search ----
    def search(self):
        self._ck(self._lib.uvg_search(self._h), "uvg_search")

    def search_async(self):
        self._ck(self._lib.uvg_search_async(self._h), "uvg_search_async")

    def wait(self):
        self._ck(self._lib.uvg_wait(self._h), "uvg_wait")

    def reset_search(self):
        self._ck(self._lib.uvg_reset_search(self._h), "uvg_reset_search")

    def heap(self, iq):
        """-> (scores int32[n,6], refs int64[n]) in the reference's output order"""
        buf = (_capi.uvg_item * self.K)()
        n = C.c_int32(0)
        self._ck(self._lib.uvg_get_heap(self._h, int(iq), buf, C.byref(n)), "uvg_get_heap")
        arr = np.frombuffer(buf, dtype=np.dtype([("score", np.int32, 6), ("ref", np.int64)]), count=n.value)
        return arr["score"].copy(), arr["ref"].copy()

    ITEM_DTYPE = np.dtype([("score", np.int32, 6), ("ref", np.int64)])          # uvg_item

    def heaps_raw(self, out=None, n_out=None):
        """every query's table in one call (uvg_get_all_heaps): (items [n_query, max(2, nbest)] of ITEM_DTYPE, rows per
        query int32 [n_query]); rows beyond a query's count are unspecified.  `out` / `n_out`: buffers of that shape to fill
        (a serving loop that keeps its result buffers pays no page faults for fresh ones)"""
        if out is None:
            out = np.empty((self.n_query, self.K), self.ITEM_DTYPE)
        if n_out is None:
            n_out = np.empty(self.n_query, np.int32)
        assert out.dtype == self.ITEM_DTYPE and out.shape == (self.n_query, self.K) and out.flags.c_contiguous
        assert n_out.dtype == np.int32 and n_out.shape == (self.n_query,) and n_out.flags.c_contiguous
        self._ck(self._lib.uvg_get_all_heaps(self._h, out.ctypes.data, n_out.ctypes.data), "uvg_get_all_heaps")
        return out, n_out

    def heaps(self):
        """every query's table: [(scores int32[n,6], refs int64[n])] (views into the one buffer of heaps_raw)"""
        buf, n = self.heaps_raw()
        sc, rf = buf["score"], buf["ref"]                        # views into one buffer (no per-query copies)
        return [(sc[q, :k], rf[q, :k]) for q, k in enumerate(n.tolist())]

    def ever_best(self, first, n):
        flags = np.zeros(int(n), np.uint8)
        if n:
            self._ck(self._lib.uvg_get_ever_best(self._h, int(first), int(n), flags.ctypes.data), "uvg_get_ever_best")
        return flags

    def stats(self):
        s = _capi.uvg_stats()
        self._ck(self._lib.uvg_get_stats(self._h, C.byref(s)), "uvg_get_stats")
        return {k: getattr(s, k) for k, _ in _capi.uvg_stats._fields_}

    # ---- sharding ----
    @staticmethod
    def shard_of(ordinal, world, first_round=0, max_round=0, total=-1):
        r = C.c_int32(0)
        l = C.c_int64(0)
        rc = _capi.lib().uvg_shard_of(int(ordinal), int(total), int(world), int(first_round), int(max_round), C.byref(r), C.byref(l))
        if rc != 0:
            raise UvgError("uvg_shard_of: bad arguments")
        return r.value, l.value

    @staticmethod
    def nccl_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        rc = _capi.lib().uvg_nccl_unique_id(buf)
        if rc != 0:
            raise UvgError(f"uvg_nccl_unique_id failed: {_capi.lib().uvg_last_error(None).decode()}")
        return buf.raw

    def nccl_init(self, uid: bytes):
        buf = C.create_string_buffer(uid, 128)
        self._ck(self._lib.uvg_nccl_init(self._h, buf), "uvg_nccl_init")

    def set_global_count(self, n):
        self._ck(self._lib.uvg_set_global_count(self._h, int(n)), "uvg_set_global_count")

    def set_search_limit(self, n):
        """pipelined sharded hosts: search only the rounds that end at or below accepted ordinal n (-1: no limit)"""
        self._ck(self._lib.uvg_set_search_limit(self._h, int(n)), "uvg_set_search_limit")

    # ---- utilities ----
    def device_alloc(self, nbytes):
        p = self._lib.uvg_device_alloc(self._h, int(nbytes))
        if not p:
            raise UvgError(f"uvg_device_alloc({nbytes}) failed: {self._lib.uvg_last_error(self._h).decode()}")
        return p

    def device_free(self, p):
        self._lib.uvg_device_free(self._h, p)

    def synth_rows_device(self, clades: np.ndarray, seed, first, n, d_ptr, n_frac_mean=0.197, ambiguous_frac=0.0):
        clades = np.ascontiguousarray(clades, dtype=np.uint8)
        self._ck(self._lib.uvg_synth_rows_device(self._h, clades.ctypes.data, clades.shape[0], int(seed), int(first), int(n),
                                                 float(n_frac_mean), float(ambiguous_frac), d_ptr), "uvg_synth_rows_device")

    def d2h(self, d_ptr, nbytes):
        out = np.empty(int(nbytes), np.uint8)
        self._ck(self._lib.uvg_memcpy_d2h(self._h, out.ctypes.data, d_ptr, int(nbytes)), "uvg_memcpy_d2h")
        return out

    def d2h_into(self, host_ptr, d_ptr, nbytes):
        self._ck(self._lib.uvg_memcpy_d2h(self._h, host_ptr, d_ptr, int(nbytes)), "uvg_memcpy_d2h")

    def debug_pair_counts(self, first, n):
        out = np.empty((self.n_query, int(n), 4), np.int32)
        self._ck(self._lib.uvg_debug_pair_counts(self._h, int(first), int(n), out.ctypes.data), "uvg_debug_pair_counts")
        return out

    def debug_pair_counts_rows(self, q_first, q_count, first, n):
        out = np.empty((int(q_count), int(n), 4), np.int32)
        self._ck(self._lib.uvg_debug_pair_counts_rows(self._h, int(q_first), int(q_count), int(first), int(n), out.ctypes.data),
                 "uvg_debug_pair_counts_rows")
        return out

    def debug_ref_meta(self, first, n):
        """-> (non_n, c0, ci) int32[n]"""
        nn, c0, ci = (np.empty(int(n), np.int32) for _ in range(3))
        self._ck(self._lib.uvg_debug_ref_meta(self._h, int(first), int(n), nn.ctypes.data, c0.ctypes.data, ci.ctypes.data),
                 "uvg_debug_ref_meta")
        return nn, c0, ci

    def debug_unpack_ref(self, local):
        out = np.empty(self.nchar, np.uint8)
        self._ck(self._lib.uvg_debug_unpack_ref(self._h, int(local), out.ctypes.data), "uvg_debug_unpack_ref")
        return out
