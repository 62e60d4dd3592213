"""ctypes wrapper around oracle/libfa_oracle.so (the CPU restatement of the reference).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.
Parity is unpinned by the reference (it has no tests and cannot run here); see the header
of fa_oracle.cpp for the pins used instead.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "libfa_oracle.so")
    src = os.path.join(_HERE, "fa_oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B", "libfa_oracle.so"], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(build())
        L.fao_run.restype = ctypes.c_void_p
        L.fao_run.argtypes = [ctypes.c_char_p, ctypes.c_int64, ctypes.c_char_p, ctypes.c_int64,
                              ctypes.c_double, ctypes.c_int, ctypes.c_int]
        L.fao_run_csr.restype = ctypes.c_void_p
        L.fao_run_csr.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p,
                                  ctypes.c_int64, ctypes.c_double, ctypes.c_int, ctypes.c_int]
        L.fao_rules_digest.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.fao_free.argtypes = [ctypes.c_void_p]
        L.fao_scalar.restype = ctypes.c_int64
        L.fao_scalar.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.fao_time.restype = ctypes.c_double
        L.fao_time.argtypes = [ctypes.c_void_p, ctypes.c_int]
        for name in ("fao_level_size", "fao_level_groups", "fao_level_cands"):
            f = getattr(L, name)
            f.restype = ctypes.c_int64
            f.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.fao_cut.restype = ctypes.c_int64
        L.fao_cut.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        L.fao_level.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        L.fao_rules_max_len.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.fao_rules.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                ctypes.c_void_p, ctypes.c_void_p]
        L.fao_text.restype = ctypes.c_void_p
        L.fao_text.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int64)]
        L.fao_level1_counts.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        _LIB = L
    return _LIB


@dataclass
class OracleResult:
    status: int                 # 1 = the reference would die with empty.min (no itemset of size >= 2)
    n_lines: int
    min_count: int
    total_count: int
    level1_ties: bool
    number_format: bool
    freq_items: list            # rank -> item string
    level1_counts: np.ndarray
    levels: dict                # k -> (sets[n,k] int32 lexicographically sorted, counts[n] int32)
    n_groups: dict
    n_cands: dict
    raw_rules: tuple            # (ant[n,stride] padded -1, cons[n], conf[n] float64), unsorted
    rules: tuple                # kept rules in the reference's sorted order
    before_cut: dict
    after_cut: dict
    recommends: list
    freqitemset_text: bytes
    recommends_text: bytes
    times: dict = field(default_factory=dict)
    rules_digest: tuple = ()    # (n kept, multiset sum, multiset xor, order sum) as Python ints
    n_raw: int = 0              # rule counts (the rule arrays are empty after a `light` run)
    n_kept: int = 0

    def itemset_dict(self):
        d = {}
        for k, (sets, counts) in self.levels.items():
            for row, c in zip(sets.tolist(), counts.tolist()):
                d[tuple(row)] = c
        return d


def _text(L, h, which) -> bytes:
    n = ctypes.c_int64(0)
    p = L.fao_text(h, which, ctypes.byref(n))
    return ctypes.string_at(p, n.value) if n.value else b""


def _flags(fast_rules, skip_rules, fast_cands, packed, no_dedup, no_text=False) -> int:
    return ((1 if fast_rules else 0) | (2 if skip_rules else 0) | (4 if fast_cands else 0) | (8 if packed else 0) |
            (16 if no_dedup else 0) | (32 if no_text else 0))


def run(d_text: bytes, u_text: bytes = b"", min_support: float = 0.092, threads: int = 0,
        fast_rules: bool = False, skip_rules: bool = False, fast_cands: bool = False,
        packed: bool = False, no_dedup: bool = False, light: bool = False) -> OracleResult:
    """Runs Main.scala's sequence (Main.scala:23-37) on the bytes of D.dat and U.dat.  `packed` / `no_dedup` are
    the two size switches of fa_oracle.cpp (same outputs); `light` skips copying the rule arrays out."""
    L = lib()
    h = L.fao_run(d_text, len(d_text), u_text, len(u_text), float(min_support), int(threads),
                  _flags(fast_rules, skip_rules, fast_cands, packed, no_dedup))
    return _collect(L, h, light)


def run_csr(d_off, d_items, u_off=None, u_items=None, min_support: float = 0.092, threads: int = 0,
            fast_rules: bool = True, skip_rules: bool = False, fast_cands: bool = True,
            packed: bool = True, no_dedup: bool = False, light: bool = True) -> OracleResult:
    """The same run on integer item ids (CSR), identical to run() on the text form of the same rows."""
    L = lib()
    d_off = np.ascontiguousarray(d_off, dtype=np.int64)
    d_items = np.ascontiguousarray(d_items, dtype=np.int32)
    if u_off is None:
        u_off, u_items = np.zeros(1, dtype=np.int64), np.zeros(0, dtype=np.int32)
    u_off = np.ascontiguousarray(u_off, dtype=np.int64)
    u_items = np.ascontiguousarray(u_items, dtype=np.int32)
    h = L.fao_run_csr(d_off.ctypes.data, d_items.ctypes.data, len(d_off) - 1, u_off.ctypes.data, u_items.ctypes.data,
                      len(u_off) - 1, float(min_support), int(threads),
                      _flags(fast_rules, skip_rules, fast_cands, packed, no_dedup, no_text=light))
    return _collect(L, h, light)


def _collect(L, h, light: bool) -> OracleResult:
    try:
        sc = lambda i: int(L.fao_scalar(h, i))
        F, maxk = sc(4), sc(5)
        l1 = np.zeros(F, dtype=np.int32)
        if F:
            L.fao_level1_counts(h, l1.ctypes.data)
        levels, n_groups, n_cands = {}, {}, {}
        for k in range(1, maxk + 1):
            n = int(L.fao_level_size(h, k))
            sets = np.zeros((n, k), dtype=np.int32)
            counts = np.zeros(n, dtype=np.int32)
            if n:
                L.fao_level(h, k, sets.ctypes.data, counts.ctypes.data)
            levels[k] = (sets, counts)
            n_groups[k] = int(L.fao_level_groups(h, k))
            n_cands[k] = int(L.fao_level_cands(h, k))

        def rules(which, n):
            stride = max(1, int(L.fao_rules_max_len(h, which)))
            if light:
                n = 0
            ant = np.full((n, stride), -1, dtype=np.int32)
            cons = np.zeros(n, dtype=np.int32)
            conf = np.zeros(n, dtype=np.float64)
            if n:
                L.fao_rules(h, which, stride, ant.ctypes.data, cons.ctypes.data, conf.ctypes.data)
            return ant, cons, conf

        n_raw, n_kept = sc(8), sc(9)
        raw = rules(0, n_raw)
        kept = rules(1, n_kept)
        dig = (ctypes.c_uint64 * 4)()
        L.fao_rules_digest(h, dig)
        before = {l: int(L.fao_cut(h, l, 0)) for l in range(2, maxk)}
        after = {l: int(L.fao_cut(h, l, 1)) for l in range(2, maxk)}
        freq_items = _text(L, h, 0).decode().split("\n")[:-1] if F else []
        rec_text = _text(L, h, 2)
        recs = rec_text.decode().split("\n")[:-1] if rec_text else []
        return OracleResult(
            status=sc(0), n_lines=sc(1), min_count=sc(2), total_count=sc(3), level1_ties=bool(sc(6)),
            number_format=bool(sc(7)), freq_items=freq_items, level1_counts=l1, levels=levels,
            n_groups=n_groups, n_cands=n_cands, raw_rules=raw, rules=kept, before_cut=before,
            after_cut=after, recommends=recs, freqitemset_text=_text(L, h, 1), recommends_text=rec_text,
            times={n: L.fao_time(h, i) for i, n in enumerate(
                ["parse", "mine", "rules", "level1", "bitmaps", "pairs", "cand", "count", "rulegen", "sort", "match"])},
            rules_digest=tuple(int(x) for x in dig), n_raw=n_raw, n_kept=n_kept)
    finally:
        L.fao_free(h)


def threads() -> int:
    return int(lib().fao_threads())
