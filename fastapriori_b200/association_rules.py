"""AssociationRules — host-side mirror of the reference's rule + recommendation operator
(src/main/scala/com/hazzacheng/AR/AssociationRules.scala:17-190): same constructor and
`run(sc, userRDD)`, returning (user index, item string) pairs.  Rule generation, the
level-wise prune, the sort and the per-user match all run in the sm_100a kernels.
"""
from __future__ import annotations

import time

import numpy as np

from .data import TokenizedLines
from .fast_apriori import FreqItemsets, _int_or_none


class NoRulesError(RuntimeError):
    """The reference dies here with UnsupportedOperationException("empty.min")
    (AssociationRules.scala:147) when no itemset of size >= 2 exists."""


class AssociationRules:
    def __init__(self, freqItemset: FreqItemsets, freqItems, itemToRank, strict: bool = False):
        if not isinstance(freqItemset, FreqItemsets):
            raise TypeError("freqItemset must be a FreqItemsets (from FastApriori.run or Utils.loadFreqItemsetWithCount)")
        self.freqItemset = freqItemset
        self.freqItems = list(freqItems)
        self.itemToRank = itemToRank
        self.strict = strict      # True: raise NoRulesError where the reference throws; False: everyone gets "0"
        self.counters = {}

    # AssociationRules.scala:23-31
    def run(self, sc, userRDD: TokenizedLines, as_arrays: bool = False):
        eng = sc.engine
        fis = self.freqItemset
        if hasattr(fis, "resident_on") and hasattr(eng, "put_levels") and not fis.resident_on(eng):
            # the constructor's freqItemset did not come from this engine's last mining run (another context, or a
            # dump read back with Utils.loadFreqItemsetWithCount): install it, then the rule stage runs as usual
            fis.generation = eng.put_levels(fis.levels)
            fis.engine = eng
        if hasattr(eng, "recommend_ids"):  # the id -> rank map of removeRedundancy is applied on the device
            rec = self.genAssociationRules(sc, userRDD.offsets, userRDD.ids, id_to_rank=self._id_to_rank(userRDD))
        else:
            offsets, ranks = self.removeRedundancy(sc, userRDD)
            rec = self.genAssociationRules(sc, offsets, ranks)
        if as_arrays:
            return rec
        base = getattr(userRDD, "index_base", 0)
        items = self.freqItems
        return [(base + i, items[r] if r >= 0 else "0") for i, r in enumerate(rec.tolist())]

    # AssociationRules.scala:33-64 — per line: keep frequent items, map to ranks.  Here every token is mapped to its
    # rank or to -1 (not frequent) and the line offsets stay as they are: K9 ignores ranks outside [0, F), duplicates in
    # a line collapse in its bit mask, and a line with no frequent item gets no recommendation ("0", :49).  Identical
    # baskets are NOT merged (the reference merges them to save work; the results are the same).
    def _id_to_rank(self, userRDD: TokenizedLines) -> np.ndarray:
        id_to_rank = np.full(max(len(userRDD.vocab), 1), -1, dtype=np.int32)
        index = userRDD.token_index() if hasattr(userRDD, "token_index") else {t: j for j, t in enumerate(userRDD.vocab)}
        for tok, r in self.itemToRank.items():   # F frequent items, not the whole vocabulary
            j = index.get(tok)
            if j is not None:
                id_to_rank[j] = r
        return id_to_rank

    def removeRedundancy(self, sc, userRDD: TokenizedLines):
        id_to_rank = self._id_to_rank(userRDD)
        ranks = np.take(id_to_rank, userRDD.ids) if len(userRDD.ids) else np.zeros(0, dtype=np.int32)
        return userRDD.offsets, ranks

    # AssociationRules.scala:66-113 (+ genRules :122-188, sort :116-120)
    def genAssociationRules(self, sc, offsets, ranks, id_to_rank=None):
        eng = sc.engine
        n_raw = eng.gen_rules()                                       # :126-145 (K6)
        if n_raw == 0:
            if self.strict:
                raise NoRulesError("empty.min: no frequent itemset of size >= 2 (AssociationRules.scala:147)")
            return np.full(len(offsets) - 1, -1, dtype=np.int32)
        t = time.time()
        n_kept = eng.prune_rules()                                    # :147-182 (K7)
        lvl = 2
        self.counters["cut"] = {}
        while True:
            before, after = eng.rule_cut(lvl)
            if before == 0:
                break
            self.counters["cut"][lvl] = (before, after)
            sc.log(f"==== Before cut level {lvl} Nums: {before}")     # :155
            sc.log(f"==== After cut level {lvl} Nums: {after}")       # :177
            lvl += 1
        sc.log(f"==== Use Time cut leaves Time: {int((time.time() - t) * 1000)}")
        item_int = np.zeros(len(self.freqItems), dtype=np.int32)
        self.counters["non_int_items"] = 0
        for r, s in enumerate(self.freqItems):
            v = _int_or_none(s)
            if v is None:
                # the reference throws NumberFormatException only if such an item meets a confidence tie (:119)
                self.counters["non_int_items"] += 1
                v = 0
            item_int[r] = v
        eng.sort_rules(item_int)                                      # :74 (K8)
        self.counters["raw"] = n_raw
        self.counters["rules"] = n_kept
        sc.log(f"==== Size association rules {n_kept}")               # :75
        if id_to_rank is not None:
            return eng.recommend_ids(offsets, ranks, id_to_rank)      # :45 + :80-106 (K9)
        return eng.recommend(offsets, ranks)                          # :80-106 (K9)
