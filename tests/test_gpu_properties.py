"""Size-independent properties at BASELINE config-2 scale (T40I10D1M-style, 1000 items, 0.5 %):
the oracle cannot finish this size, so the checks are structural (SURVEY.md §8c / task ③)."""
import numpy as np
import pytest

from fastapriori_b200 import AssociationRules, FastApriori, TokenizedLines, quest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mined(sc):
    n_rows, n_items = 200_000, 1000
    off, items = quest.generate(n_rows, 40, 10, 2000, n_items, seed=2)
    data = TokenizedLines.from_int_csr(off, items, n_items)
    fa = FastApriori(0.005)
    fis, i2r, fi = fa.run(sc, data)
    return data, fa, fis, i2r, fi


def test_levels_sorted_unique_and_above_threshold(mined):
    data, fa, fis, _, _ = mined
    minc = fa.counters["minCount"]
    assert fis.max_k >= 8
    for k, (sets, counts) in fis.levels.items():
        if k < 2 or len(counts) == 0:
            continue
        assert np.all(counts >= minc)
        assert np.all(sets[:, 1:] > sets[:, :-1]), "items inside a set ascend by rank"
        a, b = sets[:-1], sets[1:]
        ne = a != b
        first = np.argmax(ne, axis=1)
        assert np.all(ne.any(axis=1)), "no duplicate itemsets"
        rows = np.arange(len(first))
        assert np.all(a[rows, first] < b[rows, first]), "rows are in lexicographic order"


def test_downward_closure_and_antimonotone_counts(mined):
    _, _, fis, _, _ = mined
    rng = np.random.default_rng(0)
    for k in range(3, fis.max_k + 1):
        sets, counts = fis.levels[k]
        psets, pcounts = fis.levels[k - 1]
        lookup = {tuple(r): c for r, c in zip(psets.tolist(), pcounts.tolist())}
        idx = rng.choice(len(counts), size=min(len(counts), 300), replace=False)
        for i in idx.tolist():
            s = sets[i].tolist()
            for j in range(k):
                sub = tuple(s[:j] + s[j + 1:])
                assert sub in lookup, "every (k-1)-subset of a frequent set is frequent"
                assert lookup[sub] >= counts[i], "support is anti-monotone"


def test_supports_against_direct_recount(mined):
    data, _, fis, _, fi = mined
    # recount a sample of itemsets of every size straight from the CSR rows (numpy, no bitmaps)
    n_items = len(data.vocab)
    rank_of = np.full(n_items, -1, dtype=np.int64)
    for r, s in enumerate(fi):
        rank_of[int(s) - 1] = r
    row_of = np.repeat(np.arange(len(data)), np.diff(data.offsets))
    ranks = rank_of[data.ids]
    F = len(fi)
    rng = np.random.default_rng(1)
    for k in sorted(fis.levels):
        sets, counts = fis.levels[k]
        if k < 2 or len(counts) == 0:
            continue
        for i in rng.choice(len(counts), size=min(len(counts), 6), replace=False).tolist():
            present = np.ones(len(data), dtype=bool)
            for it in sets[i].tolist():
                has = np.zeros(len(data), dtype=bool)
                has[row_of[ranks == it]] = True
                present &= has
            assert int(present.sum()) == int(counts[i])
    assert F == len(fis.levels[1][1])


def test_rules_and_recommendations_are_consistent(sc, mined):
    data, _, fis, i2r, fi = mined
    uoff, uitems = quest.truncate_rows(*quest.generate(3000, 40, 10, 2000, 1000, seed=77), seed=3)
    users = TokenizedLines.from_int_csr(uoff, uitems, 1000)
    ar = AssociationRules(fis, fi, i2r)
    rec = ar.run(sc, users, as_arrays=True)
    n = ar.counters["rules"]
    ant, cons, conf = sc.engine.get_rules(1, n)
    assert np.all(conf[:-1] >= conf[1:]), "rules are ordered by confidence descending"
    assert np.all((conf > 0) & (conf <= 1))
    # brute-force first match for a sample of users against the dumped rule list
    off, ranks = ar.removeRedundancy(sc, users)
    lens = (ant >= 0).sum(axis=1)
    rng = np.random.default_rng(2)
    for u in rng.choice(len(users), size=40, replace=False).tolist():
        basket = np.zeros(len(fi) + 1, dtype=bool)
        basket[ranks[off[u]:off[u + 1]]] = True
        inside = np.where(ant >= 0, basket[np.where(ant >= 0, ant, len(fi))], True).all(axis=1)
        ok = inside & ~basket[cons] & (lens <= basket.sum())
        want = int(cons[np.argmax(ok)]) if ok.any() else -1
        assert int(rec[u]) == want
