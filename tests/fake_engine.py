"""A numpy stand-in for fastapriori_b200.engine.Engine, used ONLY by the CPU tests of the host-side
sharding logic (world_size-2 gloo).  It counts supports of the local row shard by brute force and sums
them across ranks through the same MiningContext hook the real engine's exchange callback uses.  It is
test infrastructure: the product path never falls back to it (MiningContext refuses to start without CUDA
unless a test injects an engine explicitly)."""
from __future__ import annotations

import numpy as np


class FakeEngine:
    device = 0

    def __init__(self):
        self.allreduce = lambda a: a      # set by the test to sc.allreduce_host
        self.levels = {}

    # transactions ---------------------------------------------------------------------------------
    def load_transactions(self, off, ids, n_items):
        self.off, self.ids, self.n_items = np.asarray(off), np.asarray(ids), int(n_items)

    def count_items(self):
        local = np.bincount(self.ids, minlength=self.n_items).astype(np.int64)
        return self.allreduce(local).astype(np.int32)

    def build_bitmaps(self, id_to_rank, rank_counts):
        r = np.asarray(id_to_rank)
        self.F = len(rank_counts)
        self.rows = []
        for i in range(len(self.off) - 1):
            rk = r[self.ids[self.off[i]:self.off[i + 1]]]
            self.rows.append(frozenset(int(x) for x in rk if x >= 0))
        self.levels = {1: (np.arange(self.F, dtype=np.int32).reshape(-1, 1), np.asarray(rank_counts, dtype=np.int32))}

    def _support(self, cands):
        local = np.array([sum(1 for row in self.rows if row.issuperset(c)) for c in cands], dtype=np.int64)
        return self.allreduce(local) if len(cands) else local

    # mining ---------------------------------------------------------------------------------------
    def count_pairs(self, min_count):
        cands = [(i, j) for i in range(self.F) for j in range(i + 1, self.F)]
        sup = self._support(cands)
        keep = [(c, s) for c, s in zip(cands, sup) if s >= min_count]
        self._store(2, keep)
        return len(keep)

    def level_counters(self, k):
        return self._counters.get(k, (0, 0))

    def gen_candidates(self, k, wait=True):
        prev = [tuple(r) for r in self.levels[k - 1][0].tolist()]
        pset = set(prev)
        self.cands, groups = [], 0
        for x in prev:
            ys = [y for y in range(x[-1] + 1, self.F)
                  if all(tuple(sorted(set(x) - {t} | {y})) in pset for t in x)]
            groups += bool(ys)
            self.cands += [x + (y,) for y in ys]
        self._counters = getattr(self, "_counters", {})
        self._counters[k] = (groups, len(self.cands))
        return (groups, len(self.cands)) if wait else None

    def count_candidates(self, k, min_count):
        sup = self._support(self.cands)
        keep = [(c, s) for c, s in zip(self.cands, sup) if s >= min_count]
        self._store(k, keep)
        return len(keep)

    def _store(self, k, keep):
        sets = np.array([c for c, _ in keep], dtype=np.int32).reshape(-1, k)
        counts = np.array([s for _, s in keep], dtype=np.int32)
        self.levels[k] = (sets, counts)

    def level_size(self, k):
        return len(self.levels.get(k, (None, []))[1])

    def get_level(self, k):
        return self.levels[k]

    # rules ----------------------------------------------------------------------------------------
    def gen_rules(self):
        sup = {tuple(r): int(c) for k in self.levels for r, c in zip(self.levels[k][0].tolist(), self.levels[k][1].tolist())}
        self.conf = {}
        for s, c in sup.items():
            if len(s) >= 2:
                for i in s:
                    a = tuple(x for x in s if x != i)
                    self.conf[(a, i)] = c / sup[a]
        return len(self.conf)

    def prune_rules(self):
        self.kept = {}
        for (a, c) in sorted(self.conf, key=lambda r: len(r[0])):
            cf = self.conf[(a, c)]
            if len(a) == 1 or all(((tuple(x for x in a if x != t), c) in self.kept and
                                   self.kept[(tuple(x for x in a if x != t), c)] < cf) for t in a):
                self.kept[(a, c)] = cf
        return len(self.kept)

    def rule_cut(self, level):
        before = sum(1 for (a, _) in self.conf if len(a) == level)
        return before, sum(1 for (a, _) in self.kept if len(a) == level)

    def sort_rules(self, item_int):
        self.order = sorted(self.kept, key=lambda r: (-self.kept[r], int(item_int[r[1]])))

    def recommend(self, uoff, uranks):
        out = np.full(len(uoff) - 1, -1, dtype=np.int32)
        for u in range(len(uoff) - 1):
            b = set(int(x) for x in uranks[uoff[u]:uoff[u + 1]])
            for a, c in self.order:
                if c not in b and b.issuperset(a):
                    out[u] = c
                    break
        return out

    def stats(self):
        return {}

    def close(self):
        pass
