"""ctypes face of libomnia_recall.so (include/omnia_recall.h).

There is deliberately no fallback: if the shared library is missing this raises, and if it
loads on a machine without a B200 every omr_open fails with the library's own message."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "lib" / "libomnia_recall.so"

OK, E_INVALID, E_CUDA, E_DIM, E_CAPACITY, E_NOT_FOUND, E_WORKSPACE, E_CANCELLED = 0, -1, -2, -3, -4, -5, -6, -7
ROW_FORGOTTEN, ROW_SUPERSEDED, ROW_HAS_EMBEDDING, ROW_MUTABLE = 1, 2, 4, 8
MATCH_ANY, MATCH_NULL_ONLY, MATCH_EQ, MATCH_NULL_OR_EQ = 0, 1, 2, 3
NO_EXPIRY = 2**63 - 1
MAX_DIM = 2000
NOALLOC_MAX_ROWS, NOALLOC_MAX_K = 524288, 128

c_str_p = C.POINTER(C.c_char_p)


class Config(C.Structure):
    _fields_ = [("dim", C.c_int32), ("device", C.c_int32), ("capacity", C.c_int64), ("row_base", C.c_int64),
                ("row_stride", C.c_int64), ("candidate_margin", C.c_float), ("sub_bucket", C.c_int32),
                ("rescore_width", C.c_int32), ("force_path", C.c_int32)]


class Bulk(C.Structure):
    _fields_ = [("n", C.c_int64), ("vectors", C.c_void_p), ("observation_ids", C.c_void_p), ("entity_ids", C.c_void_p),
                ("workspaces", c_str_p), ("n_workspaces", C.c_int32), ("workspace_idx", C.c_void_p),
                ("users", c_str_p), ("n_users", C.c_int32), ("user_idx", C.c_void_p),
                ("agents", c_str_p), ("n_agents", C.c_int32), ("agent_idx", C.c_void_p),
                ("kinds", c_str_p), ("n_kinds", C.c_int32), ("kind_idx", C.c_void_p),
                ("purposes", c_str_p), ("n_purposes", C.c_int32), ("purpose_idx", C.c_void_p),
                ("confidence", C.c_void_p), ("flags", C.c_void_p), ("valid_until_us", C.c_void_p)]


class Row(C.Structure):
    _fields_ = [("observation_id", C.c_uint8 * 16), ("entity_id", C.c_uint8 * 16), ("workspace_id", C.c_char_p),
                ("virtual_user_id", C.c_char_p), ("agent_id", C.c_char_p), ("kind", C.c_char_p),
                ("purpose", C.c_char_p), ("confidence", C.c_float), ("flags", C.c_uint32),
                ("valid_until_us", C.c_int64)]


class Filter(C.Structure):
    _fields_ = [("workspace_id", C.c_char_p), ("virtual_user_id", C.c_char_p), ("user_mode", C.c_int32),
                ("agent_id", C.c_char_p), ("agent_mode", C.c_int32), ("kinds", c_str_p), ("n_kinds", C.c_int32),
                ("purposes", c_str_p), ("n_purposes", C.c_int32), ("min_confidence", C.c_float),
                ("max_distance", C.c_double), ("now_us", C.c_int64)]


class Hit(C.Structure):
    _fields_ = [("observation_id", C.c_uint8 * 16), ("entity_id", C.c_uint8 * 16), ("distance", C.c_double),
                ("row", C.c_int64)]


class SelfJoinOpts(C.Structure):
    _fields_ = [("workspace_id", C.c_char_p), ("similarity_floor", C.c_float), ("limit", C.c_int32),
                ("require_same_kind", C.c_int32), ("require_mutable", C.c_int32), ("part_index", C.c_int32),
                ("part_count", C.c_int32), ("cancel", C.POINTER(C.c_int32))]


class Pair(C.Structure):
    _fields_ = [("canonical_id", C.c_uint8 * 16), ("dupe_id", C.c_uint8 * 16), ("similarity", C.c_double),
                ("canonical_row", C.c_int64), ("dupe_row", C.c_int64), ("kind", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_int64), ("searches", C.c_int64), ("queries", C.c_int64),
                ("rescued_rows", C.c_int64), ("fallback_queries", C.c_int64), ("last_path", C.c_int32),
                ("last_sub_bucket", C.c_int32), ("last_rescore_width", C.c_int32), ("num_sms", C.c_int32),
                ("scorer_ms_sum", C.c_double), ("scorer_timed", C.c_int64), ("host_calls", C.c_int64),
                ("host_batches", C.c_int64), ("listed_entries", C.c_int64),
                ("phase_ms_sum", C.c_double * 3), ("scorer_kernel_ms_sum", C.c_double),
                ("scorer_kernel_timed", C.c_int64), ("join_candidates", C.c_int64),
                ("scorer_start_skew_ms_sum", C.c_double),
                ("scorer_end_skew_ms_sum", C.c_double), ("scorer_gap_ms_sum", C.c_double),
                ("post_span_ms_sum", C.c_double), ("post_end_after_scorer_ms_sum", C.c_double)]


# name -> (restype, argtypes).  tests/test_abi.py checks this table against the header.
SIGNATURES = {
    "omr_open": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "omr_close": (None, [C.c_void_p]),
    "omr_last_error": (C.c_char_p, []),
    "omr_count": (C.c_int64, [C.c_void_p]),
    "omr_bulk_load": (C.c_int, [C.c_void_p, C.POINTER(Bulk)]),
    "omr_upsert": (C.c_int, [C.c_void_p, C.POINTER(Row), C.c_void_p, C.c_int32]),
    "omr_update_embedding": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]),
    "omr_set_row_flags": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]),
    "omr_set_entity_flags": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]),
    "omr_set_valid_until": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "omr_set_confidence": (C.c_int, [C.c_void_p, C.c_void_p, C.c_float]),
    "omr_delete_entity": (C.c_int, [C.c_void_p, C.c_void_p]),
    "omr_delete_observation": (C.c_int, [C.c_void_p, C.c_void_p]),
    "omr_delete_scope": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int64, C.POINTER(C.c_int64)]),
    "omr_rescope": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_char_p, C.c_char_p]),
    "omr_merge_entities": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]),
    "omr_clear": (C.c_int, [C.c_void_p]),
    "omr_projection_rows": (C.c_int, [C.c_void_p, C.POINTER(Filter), C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "omr_pca_moments": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "omr_pca_project": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p]),
    "omr_search": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(Filter), C.c_int32, C.c_void_p,
                             C.c_int32, C.c_void_p, C.c_void_p]),
    "omr_search_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(Filter), C.c_int32,
                                  C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "omr_search_rows_begin": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(Filter), C.c_int32,
                                        C.c_void_p, C.c_int32, C.POINTER(C.c_int32)]),
    "omr_search_rows_finish": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "omr_cosine_ranked": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(Filter), C.c_int32, C.c_void_p,
                                    C.POINTER(C.c_int32)]),
    "omr_selfjoin": (C.c_int, [C.c_void_p, C.POINTER(SelfJoinOpts), C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "omr_upsert_batch": (C.c_int, [C.c_void_p, C.POINTER(Row), C.c_void_p, C.c_int32, C.c_int32]),
    "omr_compact": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "omr_free_slots": (C.c_int64, [C.c_void_p]),
    "omr_row_entity": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(C.c_uint32)]),
    "omr_selfjoin_rows": (C.c_int, [C.c_void_p, C.POINTER(SelfJoinOpts), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                    C.POINTER(C.c_int64)]),
    "omr_bitmap_words": (C.c_int64, [C.c_void_p]),
    "omr_device_build_bitmaps": (C.c_int, [C.c_void_p, C.POINTER(Filter), C.c_int32, C.c_void_p, C.c_void_p]),
    "omr_device_search": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                    C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "omr_device_search_begin": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                          C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.POINTER(C.c_int32)]),
    "omr_device_search_finish": (C.c_int, [C.c_void_p, C.c_int32]),
    "omr_exchange_alloc": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_void_p),
                                     C.POINTER(C.c_int64)]),
    "omr_ipc_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "omr_exchange_setup": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "omr_device_search_exchange_begin": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p,
                                                   C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.POINTER(C.c_int32)]),
    "omr_device_search_exchange_finish": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "omr_device_merge_topk": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_void_p]),
    "omr_exact_topk_noalloc": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p]),
    "omr_get_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "omr_set_profiling": (C.c_int, [C.c_void_p, C.c_int32]),
    "omr_kind_name": (C.c_char_p, [C.c_void_p, C.c_uint32]),
    "omr_debug_time_scorer": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_float)]),
    "omr_debug_post_timing": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32]),
    "omr_debug_candidate_keys": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p,
                                           C.c_int64]),
}

_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(
                f"{LIB_PATH} is missing: run `python -m omnia_b200.build` (needs nvcc). "
                "omnia_b200 has no CPU or PyTorch fallback.")
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class RecallError(RuntimeError):
    """Mirrors the Go `error` the reference returns ("memory: ..." prefix, store_read.go:292)."""

    def __init__(self, code: int, message: str):
        super().__init__(message)
        self.code = code


class NotFound(RecallError):
    """memory.ErrNotFound (internal/memory/store.go:69)."""


def check(rc: int) -> None:
    if rc != OK:
        msg = (load().omr_last_error() or b"").decode("utf-8", "replace")
        raise (NotFound if rc == E_NOT_FOUND else RecallError)(rc, msg or f"omr error {rc}")
