"""Shared helpers of the parity tests: run the product path (through the C-ABI) and the oracle on
the same bytes and compare stage by stage."""
from __future__ import annotations

import numpy as np

from fastapriori_b200 import AssociationRules, FastApriori, TokenizedLines
from fastapriori_b200 import utils as Utils
from oracle import oracle_api


def run_product(sc, d_text: bytes, u_text: bytes, min_support: float):
    data = TokenizedLines.from_text(d_text)
    users = TokenizedLines.from_text(u_text)
    fa = FastApriori(min_support, 0)
    fis, item_to_rank, freq_items = fa.run(sc, data)
    ar = AssociationRules(fis, freq_items, item_to_rank)
    rec = ar.run(sc, users)
    return {"fa": fa, "ar": ar, "fis": fis, "item_to_rank": item_to_rank, "freq_items": freq_items,
            "recommends": rec, "freqitemset_text": Utils.freqItemsetLines(fis, freq_items)}


def canonical_rules(ant, cons, conf):
    """Order-insensitive view: sorted list of (antecedent tuple, consequent, confidence bits)."""
    out = []
    bits = np.asarray(conf, dtype=np.float64).view(np.uint64)
    for a, c, b in zip(ant.tolist(), cons.tolist(), bits.tolist()):
        out.append((tuple(x for x in a if x >= 0), c, b))
    return sorted(out)


def check_against_oracle(sc, d_text: bytes, u_text: bytes, min_support: float, fast_rules: bool = True):
    """Bit-exact comparison of every stage.  Returns (product, oracle) for extra assertions."""
    ora = oracle_api.run(d_text, u_text, min_support, fast_rules=fast_rules)
    got = run_product(sc, d_text, u_text, min_support)
    fa, ar, fis = got["fa"], got["ar"], got["fis"]
    # a1/a2: count, minCount, rank order, level-1 counts
    assert fa.counters["count"] == ora.n_lines
    assert fa.counters["minCount"] == ora.min_count
    assert got["freq_items"] == ora.freq_items
    assert fis.levels[1][1].tolist() == ora.level1_counts.tolist()
    # a4..a7: every level, sets and supports, plus the reference's println counters
    max_k = max([k for k in ora.levels if len(ora.levels[k][1])], default=0)
    for k in range(2, max(max_k, fis.max_k) + 1):
        o_sets, o_counts = ora.levels.get(k, (np.zeros((0, k), np.int32), np.zeros(0, np.int32)))
        g_sets, g_counts = fis.levels.get(k, (np.zeros((0, k), np.int32), np.zeros(0, np.int32)))
        assert g_sets.shape == o_sets.shape, f"level {k}: {g_sets.shape} vs {o_sets.shape}"
        assert np.array_equal(g_sets, o_sets), f"level {k} itemsets differ"
        assert np.array_equal(g_counts, o_counts), f"level {k} supports differ"
        if k >= 3 and k in fa.counters:
            assert fa.counters[k]["groups"] == ora.n_groups[k], f"level {k} group count"
            assert fa.counters[k]["candidates"] == ora.n_cands[k], f"level {k} candidate count"
    assert got["freqitemset_text"] == ora.freqitemset_text
    eng = sc.engine
    if ora.status == 1:
        # the reference throws empty.min here (A-16); the product's declared behaviour: everyone gets "0"
        assert all(r == "0" for _, r in got["recommends"])
        return got, ora
    # a8: raw rules (confidence to the last bit), prune counters, kept rules
    n_raw = ar.counters["raw"]
    assert n_raw == len(ora.raw_rules[1])
    assert canonical_rules(*eng.get_rules(0, n_raw)) == canonical_rules(*ora.raw_rules)
    for lvl, (b, a) in ar.counters["cut"].items():
        assert b == ora.before_cut.get(lvl, 0) and a == ora.after_cut.get(lvl, 0), f"cut level {lvl}"
    n_kept = ar.counters["rules"]
    assert n_kept == len(ora.rules[1])
    g_ant, g_cons, g_conf = eng.get_rules(1, n_kept)
    assert canonical_rules(g_ant, g_cons, g_conf) == canonical_rules(*ora.rules)
    # a9: order — identical (conf, consequent) sequence; order inside full ties is free (A-13)
    assert np.array_equal(g_conf.view(np.uint64), ora.rules[2].view(np.uint64))
    assert np.array_equal(g_cons, ora.rules[1])
    # a10/a11: recommendations, one per user line, in line order
    assert [r for _, r in got["recommends"]] == ora.recommends
    assert [i for i, _ in got["recommends"]] == list(range(len(ora.recommends)))
    return got, ora
