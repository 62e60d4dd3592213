"""ctypes binding of libfastapriori_b200.so (include/fastapriori_b200.h).

The library is the product: if it is missing or cannot create a context on a CUDA device the
package fails loudly.  There is no Python or CPU substitute for any entry point.
"""
from __future__ import annotations

import ctypes
import os

from . import build as _build

c_i32p = ctypes.POINTER(ctypes.c_int32)
c_i64p = ctypes.POINTER(ctypes.c_int64)
ALLREDUCE_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p)


class FaStats(ctypes.Structure):
    _fields_ = [(n, ctypes.c_double) for n in
                ("ms_hist", "ms_bitmap", "ms_pairs", "ms_candgen", "ms_count", "ms_threshold", "ms_rules",
                 "ms_sort", "ms_match", "ms_exchange", "count_bytes", "count_load_bytes", "pair_ops")] + \
               [(n, ctypes.c_int64) for n in ("count_launches", "kernel_launches", "bitmap_bytes", "row_trims", "live_rows", "total_count", "weighted",
                                              "distinct_baskets", "match_retries")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


# name -> (restype, argtypes); every symbol include/fastapriori_b200.h declares
SIGNATURES = {
    "fa_create": (ctypes.c_int, [ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]),
    "fa_destroy": (None, [ctypes.c_void_p]),
    "fa_last_error": (ctypes.c_char_p, [ctypes.c_void_p]),
    "fa_set_exchange": (ctypes.c_int, [ctypes.c_void_p, ALLREDUCE_FN, ctypes.c_void_p]),
    "fa_comm_unique_id": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int64]),
    "fa_comm_init": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32]),
    "fa_load_transactions": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32]),
    "fa_attach_transactions": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32]),
    "fa_load_transactions_u16": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32]),
    "fa_count_items": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p]),
    "fa_global_rows": (ctypes.c_int, [ctypes.c_void_p, c_i64p]),
    "fa_build_bitmaps": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p]),
    "fa_count_pairs": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, c_i64p]),
    "fa_gen_candidates": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, c_i64p, c_i64p]),
    "fa_count_candidates": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, c_i64p]),
    "fa_level_counters": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, c_i64p, c_i64p]),
    "fa_conf_key_bits": (ctypes.c_int, [ctypes.c_int64, ctypes.c_int64, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]),
    "fa_level_size": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, c_i64p]),
    "fa_get_level": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p]),
    "fa_get_level_async": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p]),
    "fa_put_level": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64]),
    "fa_wait_copies": (ctypes.c_int, [ctypes.c_void_p]),
    "fa_gen_rules": (ctypes.c_int, [ctypes.c_void_p, c_i64p]),
    "fa_prune_rules": (ctypes.c_int, [ctypes.c_void_p, c_i64p]),
    "fa_rule_cut": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, c_i64p, c_i64p]),
    "fa_sort_rules": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p]),
    "fa_rules_max_len": (ctypes.c_int, [ctypes.c_void_p, c_i32p]),
    "fa_get_rules": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "fa_rules_digest": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p]),
    "fa_recommend": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]),
    "fa_recommend_ids": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p]),
    "fa_get_stats": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(FaStats)]),
    "fa_reset_stats": (ctypes.c_int, [ctypes.c_void_p]),
    "fa_probe_l1_bandwidth": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double)]),
    "fa_pair_counts_device": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), c_i32p]),
    "fa_text_parse": (ctypes.c_int, [ctypes.c_char_p, ctypes.c_int64, ctypes.POINTER(ctypes.c_void_p)]),
    "fa_text_rows": (ctypes.c_int64, [ctypes.c_void_p]),
    "fa_text_nnz": (ctypes.c_int64, [ctypes.c_void_p]),
    "fa_text_vocab_size": (ctypes.c_int32, [ctypes.c_void_p]),
    "fa_text_vocab_bytes": (ctypes.c_int64, [ctypes.c_void_p]),
    "fa_text_copy": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "fa_text_free": (None, [ctypes.c_void_p]),
    "fa_lines_new": (ctypes.c_int, [ctypes.c_char_p, ctypes.c_int64, ctypes.c_int32, ctypes.POINTER(ctypes.c_void_p)]),
    "fa_lines_add_level": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32]),
    "fa_lines_finish": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, c_i64p]),
    "fa_lines_copy": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p]),
    "fa_lines_data": (ctypes.c_void_p, [ctypes.c_void_p]),
    "fa_lines_free": (None, [ctypes.c_void_p]),
    "fa_count_groups_noalloc": (ctypes.c_int, [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32,
                                               ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                               ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, c_i64p, ctypes.c_char_p, ctypes.c_int64]),
}

_LIB = None


class NativeError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libfastapriori_b200 error {code}: {message}")
        self.code = code


def load(path: str | None = None):
    """Loads the shared library (building it in-tree if the sources are newer) and binds every
    declared symbol.  Raises if the library or a symbol is missing."""
    global _LIB
    if _LIB is not None and path is None:
        return _LIB
    so = path or _build.LIB
    if path is None and os.path.isdir(_build.CSRC):
        try:
            so = _build.build_cuda()
        except Exception as e:  # no nvcc on this machine: use the prebuilt file if it is there
            if not os.path.exists(so):
                raise RuntimeError(f"libfastapriori_b200.so is not built and nvcc failed: {e}") from e
    lib = ctypes.CDLL(so)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _LIB = lib
    return lib
