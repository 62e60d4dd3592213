"""Engine — one fa_ctx (one GPU) behind numpy-typed methods, 1:1 with the C-ABI entry points.

This is the only module that calls into libfastapriori_b200.so.  It holds no algorithmic
logic and no fallback: every method is a single C call plus buffer marshalling.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _native
from ._native import ALLREDUCE_FN, FaStats, NativeError


def _i32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int32)


def _i64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int64)


class Engine:
    def __init__(self, device: int = 0, stream: int | None = None, lib_path: str | None = None):
        self._lib = _native.load(lib_path)
        self._ctx = ctypes.c_void_p()
        rc = self._lib.fa_create(int(device), ctypes.c_void_p(stream or 0), ctypes.byref(self._ctx))
        if rc != 0 or not self._ctx:
            raise NativeError(rc, f"fa_create(device={device}) failed: no usable CUDA device (there is no CPU fallback)")
        self.device = int(device)
        self._cb = None
        self._keep = []
        self.n_items = 0
        self.n_frequent = 0

    # -- plumbing ---------------------------------------------------------------------------
    def _check(self, rc: int):
        if rc != 0:
            raise NativeError(rc, (self._lib.fa_last_error(self._ctx) or b"").decode())

    def close(self):
        if self._ctx:
            self._lib.fa_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_exchange(self, fn):
        """fn(dev_ptr: int, count: int, stream: int) sums `count` int32 at dev_ptr over all ranks, in place."""
        if fn is None:
            self._cb = None
            self._check(self._lib.fa_set_exchange(self._ctx, ALLREDUCE_FN(), None))
            return

        def _cb(user, dev, count, stream):
            try:
                fn(int(dev or 0), int(count), int(stream or 0))
                return 0
            except Exception as e:  # surfaces as FA_EXCHANGE on the calling entry point
                self._cb_error = e
                return 1

        self._cb = ALLREDUCE_FN(_cb)
        self._check(self._lib.fa_set_exchange(self._ctx, self._cb, None))

    @staticmethod
    def comm_unique_id() -> bytes:
        """fa_comm_unique_id: the 128 bytes one rank creates and every rank passes to comm_init."""
        lib = _native.load()
        buf = ctypes.create_string_buffer(128)
        err = ctypes.create_string_buffer(512)
        rc = lib.fa_comm_unique_id(buf, err, 512)
        if rc != 0:
            raise RuntimeError(f"fa_comm_unique_id failed ({rc}): {err.value.decode(errors='replace')}")
        return buf.raw

    def comm_init(self, unique_id: bytes, rank: int, world_size: int):
        """fa_comm_init (collective): from here on the library all-reduces its count vectors itself over NCCL."""
        if len(unique_id) != 128:
            raise ValueError("unique_id must be the 128 bytes of comm_unique_id()")
        buf = ctypes.create_string_buffer(unique_id, 128)
        self._check(self._lib.fa_comm_init(self._ctx, buf, int(rank), int(world_size)))

    # -- transactions -----------------------------------------------------------------------
    def load_transactions(self, row_offsets, item_ids, n_items: int):
        off = _i64(row_offsets)
        self.n_items = int(n_items)
        if getattr(item_ids, "dtype", None) == np.uint16:  # compact ids (TokenizedLines.compact()): half the PCIe bytes
            ids = np.ascontiguousarray(item_ids)
            self._keep = [off, ids]  # async H2D from pinned memory: keep the arrays alive
            self._check(self._lib.fa_load_transactions_u16(self._ctx, off.ctypes.data, ids.ctypes.data, len(off) - 1, int(n_items)))
            return
        ids = _i32(item_ids)
        self._keep = [off, ids]
        self._check(self._lib.fa_load_transactions(self._ctx, off.ctypes.data, ids.ctypes.data, len(off) - 1, int(n_items)))

    def attach_transactions(self, dev_offsets_ptr: int, dev_items_ptr: int, n_rows: int, n_items: int):
        self.n_items = int(n_items)
        self._check(self._lib.fa_attach_transactions(self._ctx, ctypes.c_void_p(dev_offsets_ptr),
                                                     ctypes.c_void_p(dev_items_ptr), int(n_rows), int(n_items)))

    # -- mining ------------------------------------------------------------------------------
    def count_items(self) -> np.ndarray:
        out = np.zeros(self.n_items, dtype=np.int32)
        self._check(self._lib.fa_count_items(self._ctx, out.ctypes.data))
        return out

    def global_rows(self) -> int:
        """Rows of all shards (summed along with the level-1 counts by count_items)."""
        n = ctypes.c_int64()
        self._check(self._lib.fa_global_rows(self._ctx, ctypes.byref(n)))
        return n.value

    def build_bitmaps(self, id_to_rank, rank_counts):
        m, c = _i32(id_to_rank), _i32(rank_counts)
        self.n_frequent = len(c)
        self.generation = getattr(self, "generation", 0) + 1  # a new lattice replaces whatever was resident
        self._check(self._lib.fa_build_bitmaps(self._ctx, m.ctypes.data, len(c), c.ctypes.data))

    def count_pairs(self, min_count: int) -> int:
        n = ctypes.c_int64()
        self._check(self._lib.fa_count_pairs(self._ctx, int(min_count), ctypes.byref(n)))
        return n.value

    def gen_candidates(self, k: int, wait: bool = True):
        """K5.  wait=False queues the join without a read-back (the level then costs one host/device round trip);
        the (groups, candidates) totals come from level_counters(k) after count_candidates(k)."""
        if not wait:
            self._check(self._lib.fa_gen_candidates(self._ctx, int(k), None, None))
            return None
        g, c = ctypes.c_int64(), ctypes.c_int64()
        self._check(self._lib.fa_gen_candidates(self._ctx, int(k), ctypes.byref(g), ctypes.byref(c)))
        return g.value, c.value

    def level_counters(self, k: int):
        g, c = ctypes.c_int64(), ctypes.c_int64()
        self._check(self._lib.fa_level_counters(self._ctx, int(k), ctypes.byref(g), ctypes.byref(c)))
        return g.value, c.value

    def count_candidates(self, k: int, min_count: int) -> int:
        n = ctypes.c_int64()
        self._check(self._lib.fa_count_candidates(self._ctx, int(k), int(min_count), ctypes.byref(n)))
        return n.value

    def level_size(self, k: int) -> int:
        n = ctypes.c_int64()
        self._check(self._lib.fa_level_size(self._ctx, int(k), ctypes.byref(n)))
        return n.value

    def get_level(self, k: int):
        n = self.level_size(k)
        sets = np.empty((n, k), dtype=np.int32)
        counts = np.empty(n, dtype=np.int32)
        if n:
            self._check(self._lib.fa_get_level(self._ctx, int(k), sets.ctypes.data, counts.ctypes.data))
        return sets, counts

    def get_level_async(self, k: int):
        """Enqueues the copy of level k into page-locked host arrays and returns them at once; they hold the data
        after wait_copies().  The arrays come from torch's caching pinned allocator, so after the first run no
        page-locking happens on this path."""
        import torch
        n = self.level_size(k)
        sets_t = torch.empty((n, k), dtype=torch.int32, pin_memory=n > 0)
        counts_t = torch.empty((n,), dtype=torch.int32, pin_memory=n > 0)
        sets, counts = sets_t.numpy(), counts_t.numpy()
        if n:
            self._check(self._lib.fa_get_level_async(self._ctx, int(k), sets.ctypes.data, counts.ctypes.data))
        return sets, counts

    def put_levels(self, levels: dict):
        """fa_put_level for k = 1..K: installs host-side itemsets (k -> (sets[n,k], counts[n]), rows in lexicographic
        rank order) as the resident lattice, e.g. to run the rule stage on a dump of an earlier run."""
        self.generation = getattr(self, "generation", 0) + 1
        ks = sorted(k for k in levels if len(levels[k][1]) or k == 1)
        if not ks or ks != list(range(1, len(ks) + 1)):
            raise ValueError("levels must be 1..K without gaps")
        for k in ks:
            sets, counts = levels[k]
            sets = np.ascontiguousarray(sets, dtype=np.int32).reshape(-1, k)
            counts = _i32(counts)
            if len(sets) != len(counts):
                raise ValueError(f"level {k}: {len(sets)} rows but {len(counts)} counts")
            self._check(self._lib.fa_put_level(self._ctx, k, sets.ctypes.data if k >= 2 else None, counts.ctypes.data, len(counts)))
        self.n_frequent = len(levels[1][1])
        return self.generation

    def wait_copies(self):
        self._check(self._lib.fa_wait_copies(self._ctx))

    # -- rules -------------------------------------------------------------------------------
    def gen_rules(self) -> int:
        n = ctypes.c_int64()
        self._check(self._lib.fa_gen_rules(self._ctx, ctypes.byref(n)))
        return n.value

    def prune_rules(self) -> int:
        n = ctypes.c_int64()
        self._check(self._lib.fa_prune_rules(self._ctx, ctypes.byref(n)))
        return n.value

    def rule_cut(self, level: int):
        b, a = ctypes.c_int64(), ctypes.c_int64()
        self._check(self._lib.fa_rule_cut(self._ctx, int(level), ctypes.byref(b), ctypes.byref(a)))
        return b.value, a.value

    def sort_rules(self, rank_to_item_int):
        a = _i32(rank_to_item_int)
        self._check(self._lib.fa_sort_rules(self._ctx, a.ctypes.data))

    def get_rules(self, which: int, n: int):
        ml = ctypes.c_int32()
        self._check(self._lib.fa_rules_max_len(self._ctx, ctypes.byref(ml)))
        stride = max(1, ml.value)
        ant = np.full((n, stride), -1, dtype=np.int32)
        cons = np.zeros(n, dtype=np.int32)
        conf = np.zeros(n, dtype=np.float64)
        if n:
            self._check(self._lib.fa_get_rules(self._ctx, int(which), stride, ant.ctypes.data, cons.ctypes.data, conf.ctypes.data))
        return ant, cons, conf

    def rules_digest(self) -> tuple:
        """fa_rules_digest: (n kept, multiset sum, multiset xor, order sum) of the sorted rules, as Python ints."""
        out = (ctypes.c_uint64 * 4)()
        self._check(self._lib.fa_rules_digest(self._ctx, out))
        return tuple(int(x) for x in out)

    def recommend(self, user_offsets, user_ranks) -> np.ndarray:
        off, it = _i64(user_offsets), _i32(user_ranks)
        n = len(off) - 1
        out = np.full(max(n, 0), -1, dtype=np.int32)
        if n > 0:
            self._check(self._lib.fa_recommend(self._ctx, off.ctypes.data, it.ctypes.data, n, out.ctypes.data))
        return out

    def recommend_ids(self, user_offsets, user_ids, id_to_rank) -> np.ndarray:
        """fa_recommend_ids: like recommend(), the dense id -> rank map is applied on the device."""
        off, it, m = _i64(user_offsets), _i32(user_ids), _i32(id_to_rank)
        n = len(off) - 1
        out = np.full(max(n, 0), -1, dtype=np.int32)
        if n > 0:
            self._check(self._lib.fa_recommend_ids(self._ctx, off.ctypes.data, it.ctypes.data, n, m.ctypes.data, len(m), out.ctypes.data))
        return out

    # -- accounting --------------------------------------------------------------------------
    def stats(self) -> dict:
        s = FaStats()
        self._check(self._lib.fa_get_stats(self._ctx, ctypes.byref(s)))
        return s.as_dict()

    def probe_l1_bandwidth(self) -> float:
        """fa_probe_l1_bandwidth: measured GB/s of L1-hit 16-byte loads over the whole chip (K4's bounding unit)."""
        v = ctypes.c_double()
        self._check(self._lib.fa_probe_l1_bandwidth(self._ctx, ctypes.byref(v)))
        return v.value

    def pair_counts_device(self):
        """fa_pair_counts_device: (device pointer, F) of the F x F int32 pair matrix of the last pair pass."""
        p, f = ctypes.c_void_p(), ctypes.c_int32()
        self._check(self._lib.fa_pair_counts_device(self._ctx, ctypes.byref(p), ctypes.byref(f)))
        return int(p.value or 0), f.value

    def reset_stats(self):
        self._check(self._lib.fa_reset_stats(self._ctx))
