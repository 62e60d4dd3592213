"""CPU tests of the oracle itself: the hand-derived Appendix-B vector and the brute-force checker.
(The reference has no golden vectors of its own — SURVEY.md §4 — so these are the pins.)"""
import os

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from oracle import brute, oracle_api


def _read(golden_dir, name):
    with open(os.path.join(golden_dir, "appendix_b", name), "rb") as f:
        return f.read()


@pytest.mark.parametrize("fast", [False, True])
def test_appendix_b_known_answer(golden_dir, fast):
    d, u = _read(golden_dir, "D.dat"), _read(golden_dir, "U.dat")
    r = oracle_api.run(d, u, 0.092, fast_rules=fast)
    assert (r.n_lines, r.min_count, r.total_count) == (24, 3, 3)
    assert r.freq_items == ["10", "30", "20"]
    assert r.level1_counts.tolist() == [21, 14, 13]
    assert r.freqitemset_text == _read(golden_dir, "freqItemset.part-00000")
    assert r.recommends_text == _read(golden_dir, "recommends.part-00000")
    assert r.itemset_dict() == {(0,): 21, (1,): 14, (2,): 13, (0, 1): 12, (0, 2): 12, (1, 2): 6, (0, 1, 2): 6}
    assert (r.n_groups[3], r.n_cands[3]) == (1, 1)
    assert len(r.raw_rules[1]) == 9 and r.before_cut[2] == 3 and r.after_cut[2] == 1
    want = [l.split("|") for l in _read(golden_dir, "rules.txt").decode().splitlines() if not l.startswith("#")]
    ant, cons, conf = r.rules
    got = [[" ".join(sorted((r.freq_items[x] for x in a if x >= 0), key=int)), r.freq_items[c], float(f).hex()]
           for a, c, f in zip(ant.tolist(), cons.tolist(), conf.tolist())]
    want = [[" ".join(sorted(a.split(), key=int)), c, float.fromhex(f).hex()] for a, c, f in want]
    assert got == want


def test_counts_dump_matches_golden(golden_dir):
    r = oracle_api.run(_read(golden_dir, "D.dat"), b"", 0.092, skip_rules=True)
    lines = []
    for (s, c) in r.itemset_dict().items():
        lines.append(" ".join(r.freq_items[x] for x in sorted(s, reverse=True)) + f"[{c}]")
    assert sorted(lines) == _read(golden_dir, "counts.txt").decode().splitlines()


def test_tokenisation_rules():
    # \r\n, \r and \n terminate records; trim strips chars <= 0x20; a blank line is the token "" (A-18)
    d = b"  1 2\t3 \r\n1  2\r\x01 1 \n\n1 2\n"
    r = oracle_api.run(d, b"", 0.5, skip_rules=True)
    assert r.n_lines == 5
    assert r.min_count == 3
    assert r.freq_items[0] == "1" and dict(zip(r.freq_items, r.level1_counts.tolist())) == {"1": 4, "2": 3}
    b = brute.mine(d.decode(), "", 0.5)
    assert b["freq_items"] == r.freq_items


def test_occurrence_vs_presence_counts():
    # level 1 counts occurrences (A-2): "7 7" counts 7 twice; pairs use the de-duplicated row
    d = b"7 7 8\n7 8\n8\n9\n"
    r = oracle_api.run(d, b"", 0.5, skip_rules=True)
    assert dict(zip(r.freq_items, r.level1_counts.tolist())) == {"7": 3, "8": 3}
    assert r.itemset_dict()[(0, 1)] == 2
    assert r.level1_ties


def test_no_pairs_is_the_reference_crash_case():
    r = oracle_api.run(b"1\n2\n1\n2\n", b"1\n", 0.5)
    assert r.status == 1 and r.recommends == []          # AssociationRules.scala:147 empty.min
    assert r.freqitemset_text == b"1\n2\n"                # freqItemset is written before the crash


def test_fast_rules_flag_is_equivalent():
    rng = np.random.default_rng(7)
    rows = [" ".join(map(str, sorted(set(rng.integers(1, 9, size=rng.integers(1, 7)).tolist())))) for _ in range(60)]
    d = ("\n".join(rows) + "\n").encode()
    u = ("\n".join(rows[:20]) + "\n").encode()
    a = oracle_api.run(d, u, 0.1, fast_rules=False)
    b = oracle_api.run(d, u, 0.1, fast_rules=True)
    assert a.freqitemset_text == b.freqitemset_text and a.recommends == b.recommends
    assert np.array_equal(a.rules[2], b.rules[2]) and np.array_equal(a.rules[1], b.rules[1])


rows_st = st.lists(st.lists(st.integers(1, 9), min_size=0, max_size=6), min_size=1, max_size=40)


@settings(max_examples=60, deadline=None)
@given(rows=rows_st, users=rows_st, sup=st.sampled_from([0.05, 0.1, 0.2, 0.34, 0.5]))
def test_oracle_matches_brute_force(rows, users, sup):
    d = "\n".join(" ".join(map(str, r)) for r in rows) + "\n"
    u = "\n".join(" ".join(map(str, r)) for r in users) + "\n"
    o = oracle_api.run(d.encode(), u.encode(), sup)
    b = brute.mine(d, u, sup)
    assert o.n_lines == b["n"] and o.min_count == b["min_count"]
    assert o.freq_items == b["freq_items"]
    assert o.itemset_dict() == b["itemsets"]
    assert o.freqitemset_text == b["freqitemset_text"]
    if b["recommends"] is None:
        assert o.status == 1
        return
    ant, cons, conf = o.rules
    got = {(tuple(x for x in a if x >= 0), c): f for a, c, f in zip(ant.tolist(), cons.tolist(), conf.tolist())}
    assert got == b["kept"]
    assert o.recommends == b["recommends"]


@pytest.mark.parametrize("name", ["quest_t8", "quest_t12_deep"])
def test_oracle_reproduces_committed_golden(golden_dir, name):
    d = os.path.join(golden_dir, name)
    rd = lambda f: open(os.path.join(d, f), "rb").read()
    sup = float(rd("min_support.txt"))
    r = oracle_api.run(rd("D.dat"), rd("U.dat"), sup)
    assert r.freqitemset_text == rd("freqItemset.part-00000")
    assert r.recommends_text == rd("recommends.part-00000")
    counts = sorted(" ".join(r.freq_items[x] for x in sorted(s, reverse=True)) + f"[{c}]" for s, c in r.itemset_dict().items())
    assert ("\n".join(counts) + "\n").encode() == rd("counts.txt")
    fast = oracle_api.run(rd("D.dat"), rd("U.dat"), sup, fast_rules=True, fast_cands=True)
    assert fast.freqitemset_text == r.freqitemset_text and fast.recommends_text == r.recommends_text


# ---- the two size switches of the oracle (packed bitmaps, no row dedup) and the CSR entry are held to the naive mode ----
def _same_outputs(a, b, total_count=True):
    assert (a.status, a.n_lines, a.min_count) == (b.status, b.n_lines, b.min_count)
    if total_count:
        assert a.total_count == b.total_count
    assert a.freq_items == b.freq_items and a.level1_counts.tolist() == b.level1_counts.tolist()
    assert a.itemset_dict() == b.itemset_dict()
    assert a.n_groups == b.n_groups and a.n_cands == b.n_cands
    assert a.freqitemset_text == b.freqitemset_text and a.recommends_text == b.recommends_text
    assert a.before_cut == b.before_cut and a.after_cut == b.after_cut
    assert a.rules_digest == b.rules_digest and (a.n_raw, a.n_kept) == (b.n_raw, b.n_kept)
    for x, y in zip(a.rules, b.rules):
        assert np.array_equal(x.view(np.uint64) if x.dtype == np.float64 else x, y.view(np.uint64) if y.dtype == np.float64 else y)


@settings(max_examples=60, deadline=None)
@given(rows=rows_st, users=rows_st, sup=st.sampled_from([0.05, 0.1, 0.2, 0.34, 0.5]))
def test_packed_and_no_dedup_modes_match_naive(rows, users, sup):
    d = ("\n".join(" ".join(map(str, r)) for r in rows) + "\n").encode()
    u = ("\n".join(" ".join(map(str, r)) for r in users) + "\n").encode()
    naive = oracle_api.run(d, u, sup)
    _same_outputs(oracle_api.run(d, u, sup, packed=True), naive)
    _same_outputs(oracle_api.run(d, u, sup, packed=True, no_dedup=True, fast_rules=True, fast_cands=True), naive, total_count=False)
    _same_outputs(oracle_api.run(d, u, sup, no_dedup=True), naive, total_count=False)


@pytest.mark.parametrize("name", ["appendix_b", "quest_t8", "quest_t12_deep"])
def test_packed_mode_on_golden(golden_dir, name):
    d = os.path.join(golden_dir, name)
    rd = lambda f: open(os.path.join(d, f), "rb").read()
    sup = float(rd("min_support.txt")) if name != "appendix_b" else 0.092
    naive = oracle_api.run(rd("D.dat"), rd("U.dat"), sup)
    assert naive.freqitemset_text == rd("freqItemset.part-00000")
    _same_outputs(oracle_api.run(rd("D.dat"), rd("U.dat"), sup, packed=True, fast_rules=True, fast_cands=True), naive)
    _same_outputs(oracle_api.run(rd("D.dat"), rd("U.dat"), sup, packed=True, no_dedup=True), naive, total_count=False)


def test_packed_mode_heavy_and_light_words():
    # 200 copies of one basket (a heavy row, weight 200) next to 300 mostly distinct light rows: both summation paths of
    # the packed counter (per-bit weights below heavy_words, popcount above) run and must agree with the naive mode
    rng = np.random.default_rng(3)
    rows = ["1 2 3 4"] * 200 + [" ".join(map(str, sorted(set(rng.integers(1, 40, size=8).tolist())))) for _ in range(300)]
    rng.shuffle(rows)
    d = ("\n".join(rows) + "\n").encode()
    naive = oracle_api.run(d, d[:2000], 0.02)
    assert naive.total_count < naive.n_lines - 150
    _same_outputs(oracle_api.run(d, d[:2000], 0.02, packed=True), naive)


def test_csr_entry_equals_text_entry():
    from fastapriori_b200 import quest
    off, items = quest.generate(4000, 8, 3, 30, 120, seed=21)
    uoff, uitems = quest.truncate_rows(*quest.generate(300, 8, 3, 30, 120, seed=22), seed=5)
    a = oracle_api.run(quest.to_text(off, items), quest.to_text(uoff, uitems), 0.03, fast_rules=True)
    b = oracle_api.run_csr(off, items, uoff, uitems, 0.03, light=False)
    _same_outputs(b, a)
    c = oracle_api.run_csr(off, items, uoff, uitems, 0.03, light=True, no_dedup=True)
    assert c.rules_digest == a.rules_digest and c.recommends == a.recommends and c.itemset_dict() == a.itemset_dict()
