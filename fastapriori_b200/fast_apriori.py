"""FastApriori — host-side mirror of the reference's mining operator
(src/main/scala/com/hazzacheng/AR/FastApriori.scala:18-248): same constructor, setters and
`run(sc, data)` shape, same result triple.  The host decides minCount and the rank order and
drives the level loop; every pass over transactions, bitmaps or candidates is one C-ABI call
into the sm_100a kernels (fastapriori_b200.engine).
"""
from __future__ import annotations

import functools
import math
import time

import numpy as np

from .data import TokenizedLines


@functools.lru_cache(maxsize=1 << 20)  # the same few thousand item strings are parsed on every run
def _int_or_none(s: str):
    """Scala's String.toInt: optional sign, ASCII digits, Int32 range."""
    t = s[1:] if s[:1] in ("+", "-") else s
    if not t or not t.isascii() or not t.isdigit():
        return None
    v = int(s)
    return v if -2**31 <= v <= 2**31 - 1 else None


def rank_order(vocab, counts: np.ndarray, min_count: int, tie_order: str = "canonical", num_partitions: int = 500) -> np.ndarray:
    """Dense ids of frequent items in rank order (FastApriori.scala:57-62): count >= minCount,
    sorted by count descending.  Among equal counts the reference inherits Spark's collect()
    order; the canonical rule used here (and by the oracle) is: items
    that parse as Int ascending by value, then the rest in byte order."""
    ids = np.flatnonzero(counts >= min_count)
    if len(ids) == 0:
        return ids.astype(np.int32)
    if tie_order == "spark":  # the emulated collect() order, then the reference's stable sort by count
        pos = spark_collect_order(vocab, num_partitions)
        return ids[np.lexsort((pos[ids], -counts[ids].astype(np.int64)))].astype(np.int32)
    if tie_order != "canonical":
        raise ValueError(f"tie_order must be 'canonical' or 'spark', not {tie_order!r}")
    ints = [_int_or_none(vocab[i]) for i in ids]
    if all(v is not None for v in ints):
        order = np.lexsort((np.array(ints, dtype=np.int64), -counts[ids].astype(np.int64)))
        return ids[order].astype(np.int32)
    keyed = sorted(range(len(ids)), key=lambda j: (-int(counts[ids[j]]), 0 if ints[j] is not None else 1,
                                                    ints[j] if ints[j] is not None else 0, vocab[ids[j]].encode("utf-8", "surrogateescape")))
    return ids[keyed].astype(np.int32)


# ---- Spark's collect() order for rank ties ---------------------------------------------
# The reference ranks items by a STABLE sort on count (FastApriori.scala:60) of `reduceByKey(partitioner).collect()`
# (:55-58), so equal-count items keep Spark's order: partitions 0..P-1 of HashPartitioner(P), and inside a partition
# the slot order of the reduce side's AppendOnlyMap.  Restated from spark-core_2.10 2.1.0 (un-vendored dependency,
# pom.xml:26-30; not under /root/reference and no JVM here, so this emulation is unverifiable on this box):
# partition = nonNegativeMod(String.hashCode, P); AppendOnlyMap: capacity 64, doubles when size > 0.7 * capacity,
# slot = murmur3_32(hashCode) & mask, then triangular probing (+1, +2, ...).  What it cannot know is the arrival
# order of the shuffle records, which decides who keeps a contested slot: keys are inserted in order of first
# appearance in the input.  Only the position of equal-count items inside a freqItemset line depends on any of this.
def java_string_hash(s: str) -> int:
    """java.lang.String.hashCode over UTF-16 code units, as a signed 32-bit int."""
    units = s.encode("utf-16-be", "surrogatepass")
    h = 0
    for i in range(0, len(units), 2):
        h = (31 * h + ((units[i] << 8) | units[i + 1])) & 0xFFFFFFFF
    return h - (1 << 32) if h & 0x80000000 else h


def murmur3_int(h: int) -> int:
    """Guava Hashing.murmur3_32().hashInt(h).asInt() (seed 0), Spark's AppendOnlyMap.rehash."""
    k = (h & 0xFFFFFFFF) * 0xCC9E2D51 & 0xFFFFFFFF
    k = ((k << 15) | (k >> 17)) & 0xFFFFFFFF
    k = k * 0x1B873593 & 0xFFFFFFFF
    x = k
    x = ((x << 13) | (x >> 19)) & 0xFFFFFFFF
    x = (x * 5 + 0xE6546B64) & 0xFFFFFFFF
    x ^= 4
    x ^= x >> 16
    x = x * 0x85EBCA6B & 0xFFFFFFFF
    x ^= x >> 13
    x = x * 0xC2B2AE35 & 0xFFFFFFFF
    x ^= x >> 16
    return x - (1 << 32) if x & 0x80000000 else x


def spark_collect_order(vocab, num_partitions: int) -> np.ndarray:
    """Position of every vocab entry in the emulated `reduceByKey(HashPartitioner(P)).collect()` sequence."""
    P = int(num_partitions)
    parts = {}
    for i, tok in enumerate(vocab):
        h = java_string_hash(tok)
        parts.setdefault(h % P, []).append((i, murmur3_int(h)))  # Python's % is already non-negative
    pos = np.zeros(len(vocab), dtype=np.int64)
    nxt = 0
    for p in sorted(parts):
        cap, table, size = 64, {}, 0
        for i, mh in parts[p]:
            def put(tab, capacity, entry):
                slot, delta = entry[1] & (capacity - 1), 1
                while slot in tab:
                    slot = (slot + delta) & (capacity - 1)
                    delta += 1
                tab[slot] = entry
            put(table, cap, (i, mh))
            size += 1
            if size > int(0.7 * cap):  # growTable: re-insert in old slot order
                old = [table[s] for s in sorted(table)]
                cap, table = cap * 2, {}
                for e in old:
                    put(table, cap, e)
        for s in sorted(table):
            pos[table[s][0]] = nxt
            nxt += 1
    return pos


class FreqItemsets:
    """The reference's Array[(Set[Int], Int)] (FastApriori.scala:41): all frequent k-itemsets,
    k >= 2 first and the 1-itemsets last, over ranks.  Kept as one int32 matrix per level
    (rows in lexicographic rank order) instead of boxed sets; iterating yields
    (frozenset, count) pairs in the reference's order.  The device-resident copy stays in the
    engine so AssociationRules can use it without a round trip."""

    def __init__(self, levels: dict, engine=None, pending: bool = False):
        self._levels = levels  # k -> (sets[n,k], counts[n])
        self.engine = engine
        self._pending = pending  # device->host copies of the levels may still be in flight
        # which lattice of the engine this is: AssociationRules re-installs the host copy (fa_put_level) when the
        # engine has moved on to another one, or when the itemsets never were on a device (loaded from a dump)
        self.generation = getattr(engine, "generation", None) if engine is not None else None

    def resident_on(self, engine) -> bool:
        return engine is self.engine and self.generation is not None and self.generation == getattr(engine, "generation", None)

    @property
    def levels(self):
        if self._pending:
            if getattr(self.engine, "_ctx", None):   # a closed engine has finished its copies (fa_destroy waits for them)
                self.engine.wait_copies()
            self._pending = False
        return self._levels

    def __len__(self):
        return sum(len(c) for _, c in self.levels.values())

    def __iter__(self):
        ks = sorted(self.levels)
        for k in [k for k in ks if k >= 2] + [k for k in ks if k == 1]:
            sets, counts = self.levels[k]
            for row, c in zip(sets.tolist(), counts.tolist()):
                yield frozenset(row), c

    def as_dict(self):
        return {tuple(r): c for k in self.levels for r, c in zip(self.levels[k][0].tolist(), self.levels[k][1].tolist())}

    @property
    def max_k(self):
        return max([k for k, (_, c) in self.levels.items() if len(c)], default=0)


class FastApriori:
    def __init__(self, minSupport: float, numPartitions: int = 0, tie_order: str = "canonical"):
        self.minSupport = float(minSupport)
        self.numPartitions = int(numPartitions)  # sharding is by GPU, not partition; only tie_order="spark" reads it
        # "canonical": equal-count items by Int value (the oracle's rule); "spark": the emulated collect() order of
        # HashPartitioner(numPartitions or 500) — what a real Spark run would most likely print (see above)
        self.tie_order = tie_order
        self.counters = {}                       # the reference's println counters (FastApriori.scala:107-127)

    def setMinSupport(self, minSupport: float):
        self.minSupport = float(minSupport)
        return self

    def setNumPartitions(self, numPartitions: int):
        self.numPartitions = int(numPartitions)
        return self

    # FastApriori.scala:31-44
    def run(self, sc, data: TokenizedLines, fetch: bool = True, resident=None):
        """Returns (freqItemsets, itemToRank, freqItems) like the reference.  `data` is this rank's
        shard of the transactions.  `resident=(dev_offsets_ptr, dev_ids_ptr)` mines a CSR that is
        already in device memory (bench.py's device-resident arm).  `fetch=False` leaves the
        itemsets on the device (levels are then fetched lazily by the caller)."""
        eng = sc.engine
        n_items = len(data.vocab)
        if resident is None:
            eng.load_transactions(data.offsets, data.ids, n_items)
        else:
            eng.attach_transactions(resident[0], resident[1], len(data), n_items)
        counts = eng.count_items()                                                # :55-56 (K1 + exchange)
        count = None                                                              # :38 data.count(), over all shards:
        if hasattr(eng, "global_rows") and n_items > 0:                           # summed by K1's exchange, no collective of its own
            count = int(eng.global_rows())
        if count is None:
            count = int(sc.allreduce_host(np.array([len(data)], dtype=np.int64))[0])
        minCount = int(math.ceil(self.minSupport * count))                        # :39, IEEE double
        self.counters = {"count": count, "minCount": minCount}
        freqItems, itemToRank, rank_counts = self._genFreqItems(sc, data, minCount, counts)
        levels = self._genFreqItemsets(sc, minCount, len(freqItems), fetch)
        if fetch:
            levels[1] = (np.arange(len(freqItems), dtype=np.int32).reshape(-1, 1), rank_counts)
        return FreqItemsets(levels, eng, pending=fetch and hasattr(eng, "get_level_async")), itemToRank, freqItems

    # FastApriori.scala:46-86
    def _genFreqItems(self, sc, data, minCount, counts):
        eng = sc.engine
        order = rank_order(data.vocab, counts, minCount, self.tie_order, self.numPartitions or 500)  # :57-60
        freqItems = [data.vocab[i] for i in order]
        itemToRank = {s: r for r, s in enumerate(freqItems)}         # :62
        id_to_rank = np.full(len(data.vocab), -1, dtype=np.int32)
        id_to_rank[order] = np.arange(len(order), dtype=np.int32)
        rank_counts = counts[order].astype(np.int32)
        self.id_to_rank = id_to_rank
        self.counters["ties"] = bool(len(rank_counts) > 1 and np.any(rank_counts[1:] == rank_counts[:-1]))
        eng.build_bitmaps(id_to_rank, rank_counts)                   # :66-79 + :195-210 (K2)
        return freqItems, itemToRank, rank_counts

    # FastApriori.scala:88-130
    def _genFreqItemsets(self, sc, minCount, F, fetch):
        eng = sc.engine
        levels = {}
        t = time.time()
        n_k = eng.count_pairs(minCount)                              # :104 (K3 + exchange)
        self.counters[2] = {"candidates": F * (F - 1) // 2, "groups": F * (F - 1) // 2, "frequent": n_k}
        sc.log(f"==== 2 candidates items {F * (F - 1) // 2}")         # :226
        sc.log(f"==== 2 freq items {n_k}")                            # :107
        sc.log(f"==== Use Time 2 items {int((time.time() - t) * 1000)}")
        total = n_k
        fetch_level = getattr(eng, "get_level_async", eng.get_level)  # async: the copy overlaps the next levels
        # The copy of level k-1 is started between K5 and K4 of level k: it then overlaps the long counting kernel
        # instead of sharing the PCIe link with the size read-backs K5 waits on.
        # (starting it before K5 instead was measured: K5 waits 1 ms longer per step for its size read-backs, no gain)
        k = 3
        while n_k >= k:                                              # :110-111
            t = time.time()
            eng.gen_candidates(k, wait=False)                        # :113 (K5), queued: its totals come with K4's
            if fetch:
                levels[k - 1] = fetch_level(k - 1)
            n_k = eng.count_candidates(k, minCount)                  # :115 (K4 + exchange)
            groups, cands = eng.level_counters(k)
            sc.log(f"==== {k} candidate items {groups}")              # :114 prints the group count
            self.counters[k] = {"candidates": cands, "groups": groups, "frequent": n_k}
            sc.log(f"==== {k} freq items {n_k}")                      # :118
            sc.log(f"==== Use Time {k} items {int((time.time() - t) * 1000)}")
            total += n_k
            k += 1
        if fetch:
            levels[k - 1] = fetch_level(k - 1)                       # the last level built (possibly empty)
        self.counters["max_k"] = k - 1
        sc.log(f"==== Total freq items sets {total}")                 # :127
        return levels
