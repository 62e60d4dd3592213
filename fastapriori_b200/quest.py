"""Seeded IBM-Quest-style transaction generator (csrc/quest_gen.cpp) — workload tooling for
tests and bench.py.  Rows [row_begin, row_end) of the virtual dataset are identical no matter
how many ranks generate them."""
from __future__ import annotations

import ctypes

import numpy as np

from . import build as _build

_L = None


def _lib():
    global _L
    if _L is None:
        L = ctypes.CDLL(_build.build_quest())
        L.quest_generate.restype = ctypes.c_int
        L.quest_generate.argtypes = [ctypes.c_int64, ctypes.c_double, ctypes.c_double, ctypes.c_int32, ctypes.c_int32,
                                     ctypes.c_double, ctypes.c_uint64, ctypes.c_int64, ctypes.c_int64,
                                     ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_void_p)]
        L.quest_free.argtypes = [ctypes.c_void_p]
        L.quest_write_text.restype = ctypes.c_int
        L.quest_write_text.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64]
        _L = L
    return _L


def generate(n_rows: int, avg_len: float, avg_pat: float, n_pat: int, n_items: int, seed: int,
             row_begin: int = 0, row_end: int | None = None, corr: float = 0.5):
    """Returns (offsets int64[n+1], items int32[nnz]) with item ids in 1..n_items."""
    L = _lib()
    row_end = n_rows if row_end is None else row_end
    po, pi = ctypes.c_void_p(), ctypes.c_void_p()
    rc = L.quest_generate(n_rows, avg_len, avg_pat, n_pat, n_items, corr, seed, row_begin, row_end,
                          ctypes.byref(po), ctypes.byref(pi))
    if rc != 0:
        raise ValueError(f"quest_generate failed with code {rc}")
    try:
        n = row_end - row_begin
        off = np.ctypeslib.as_array(ctypes.cast(po, ctypes.POINTER(ctypes.c_int64)), (n + 1,)).copy()
        nnz = int(off[-1])
        items = (np.ctypeslib.as_array(ctypes.cast(pi, ctypes.POINTER(ctypes.c_int32)), (max(nnz, 1),))[:nnz]).copy()
    finally:
        L.quest_free(po)
        L.quest_free(pi)
    return off, items


def truncate_rows(offsets, items, seed: int):
    """U.dat-style baskets: every row cut to a random non-empty prefix."""
    rng = np.random.default_rng(seed)
    lens = np.diff(offsets)
    keep = np.minimum(lens, 1 + (rng.random(len(lens)) * lens).astype(np.int64))
    new_off = np.zeros(len(offsets), dtype=np.int64)
    np.cumsum(keep, out=new_off[1:])
    idx = np.repeat(offsets[:-1] - new_off[:-1], keep) + np.arange(int(new_off[-1]), dtype=np.int64)
    return new_off, items[idx]


def write_text(path: str, offsets, items):
    off = np.ascontiguousarray(offsets, dtype=np.int64)
    it = np.ascontiguousarray(items, dtype=np.int32)
    rc = _lib().quest_write_text(path.encode(), off.ctypes.data, it.ctypes.data, len(off) - 1)
    if rc != 0:
        raise OSError(f"cannot write {path}")


def to_text(offsets, items) -> bytes:
    return ("\n".join(" ".join(map(str, items[offsets[i]:offsets[i + 1]])) for i in range(len(offsets) - 1)) + "\n").encode()
