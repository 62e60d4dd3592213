"""SURVEY.md 8 f-4 on the GPU: identical filtered baskets merged into weighted rows (FastApriori.scala:66-79, bit-sliced
weight planes) and identical user baskets matched once (AssociationRules.scala:51-58).  Both are exact: every stage must
equal the oracle, which keeps the reference's weighted distinct rows natively, and the un-merged GPU path."""
import numpy as np
import pytest

from helpers import check_against_oracle

pytestmark = pytest.mark.gpu


def _ctx(monkeypatch, **env):
    from fastapriori_b200 import MiningContext
    for k, v in env.items():
        if v is None:
            monkeypatch.delenv(k, raising=False)
        else:
            monkeypatch.setenv(k, str(v))
    return MiningContext(device=0)   # the switches are read when the context is created


def _c1_text(rows, users, seed=1):
    from fastapriori_b200 import quest
    from fastapriori_b200.workloads import WORKLOADS, gen_rows, gen_users
    wl = dict(WORKLOADS["c1"], seed=seed)
    off, items = gen_rows(wl, rows, 0, rows)
    uoff, uitems = gen_users(wl, users, 0, users)
    return quest.to_text(off, items), quest.to_text(uoff, uitems), wl["min_support"]


@pytest.mark.parametrize("k4", ["tiles", "groups"])
def test_weighted_rows_equal_oracle_and_unweighted(monkeypatch, k4):
    d, u, sup = _c1_text(40_000, 3_000)
    outs = {}
    for mode in ("1", "0"):
        sc = _ctx(monkeypatch, FA_WEIGHTED=mode, FA_K4=k4)
        try:
            got, ora = check_against_oracle(sc, d, u, sup)
            st = sc.engine.stats()
            assert st["weighted"] == int(mode)
            if mode == "1":
                assert st["total_count"] == ora.total_count        # the reference's totalCount (FastApriori.scala:79)
                assert st["live_rows"] < ora.n_lines // 2            # the row axis really shrank
                assert st["row_trims"] == 0
            assert st["distinct_baskets"] > 0 and st["distinct_baskets"] <= 3_000
            outs[mode] = (got["fis"].as_dict(), [r for _, r in got["recommends"]])
            assert got["fis"].max_k >= 3
        finally:
            sc.close()
    assert outs["1"] == outs["0"]


def test_weighted_mode_is_chosen_when_it_pays(monkeypatch):
    """Unforced, rows are merged only when that removes three quarters of the row axis and at least 48 items are frequent
    (below that the pair pass is too small to repay the sort: measured); never above 64 frequent items."""
    from fastapriori_b200 import FastApriori, TokenizedLines
    sc = _ctx(monkeypatch, FA_WEIGHTED=None, FA_K4=None)
    try:
        # 56 frequent items, 1.2 M rows drawn from 3 000 distinct baskets (9 weight planes of one chunk each): merged
        rng = np.random.default_rng(4)
        baskets = [np.sort(rng.choice(56, size=rng.integers(2, 9), replace=False)) + 1 for _ in range(3000)]
        pick = rng.integers(0, 3000, size=1_200_000)
        off = np.zeros(1_200_001, dtype=np.int64)
        lens = np.array([len(b) for b in baskets])
        np.cumsum(lens[pick], out=off[1:])
        flat = np.concatenate(baskets).astype(np.int32)
        starts = np.concatenate(([0], np.cumsum(lens)[:-1]))
        items = flat[np.repeat(starts[pick] - off[:-1], lens[pick]) + np.arange(int(off[-1]))]
        fa = FastApriori(0.0005)
        fis, _, fi = fa.run(sc, TokenizedLines.from_int_csr(off, items, 56))
        st = sc.engine.stats()
        assert len(fi) == 56 and st["weighted"] == 1 and st["total_count"] == len({tuple(b) for b in baskets})
        merged = fis.as_dict()
        sc.close()
        sc = _ctx(monkeypatch, FA_WEIGHTED="0", FA_K4=None)
        fis0, _, fi0 = FastApriori(0.0005).run(sc, TokenizedLines.from_int_csr(off, items, 56))
        assert sc.engine.stats()["weighted"] == 0 and fi0 == fi and fis0.as_dict() == merged and fis0.max_k >= 3
        # 35 frequent items (configs[0]): the rows stay unmerged and the masks are not even built
        d, u, sup = _c1_text(60_000, 10)
        sc.close()
        sc = _ctx(monkeypatch, FA_WEIGHTED=None, FA_K4=None)
        got, ora = check_against_oracle(sc, d, u, sup)
        st = sc.engine.stats()
        assert st["weighted"] == 0 and st["total_count"] == -1
        # more than 64 frequent items: a basket is no single mask any more
        from fastapriori_b200 import quest
        off, items = quest.generate(20_000, 12, 5, 80, 300, seed=5)
        FastApriori(0.01).run(sc, TokenizedLines.from_int_csr(off, items, 300))
        st = sc.engine.stats()
        assert st["weighted"] == 0 and st["total_count"] == -1
    finally:
        sc.close()


def test_heavy_weights_span_many_planes(monkeypatch):
    # one basket repeated 70 000 times (weight bits up to 2^16) next to a few hundred light ones, over two chunks of plane 0
    rng = np.random.default_rng(9)
    rows = ["1 2 3"] * 70_000 + ["2 3 4"] * 5 + [" ".join(map(str, sorted(set(rng.integers(1, 12, size=5).tolist())))) for _ in range(40_000)]
    order = rng.permutation(len(rows))
    d = ("\n".join(rows[i] for i in order) + "\n").encode()
    u = ("\n".join(rows[i] for i in order[:2000]) + "\n").encode()
    for k4 in ("tiles", "groups"):
        sc = _ctx(monkeypatch, FA_WEIGHTED="1", FA_K4=k4)
        try:
            got, ora = check_against_oracle(sc, d, u, 0.0005)
            st = sc.engine.stats()
            assert st["weighted"] == 1 and st["total_count"] == ora.total_count
            assert got["fis"].as_dict()[(0, 1, 2)] >= 70_000 or max(got["fis"].as_dict().values()) >= 70_000
        finally:
            sc.close()


def test_user_baskets_are_matched_once(monkeypatch):
    d, u, sup = _c1_text(30_000, 5_000)
    recs = {}
    for dedup in ("1", "0"):
        sc = _ctx(monkeypatch, FA_WEIGHTED=None, FA_K4=None, FA_USER_DEDUP=dedup)
        try:
            got, ora = check_against_oracle(sc, d, u, sup)
            st = sc.engine.stats()
            if dedup == "1":
                ranks = {s: r for r, s in enumerate(got["freq_items"])}
                distinct = {frozenset(ranks[t] for t in line.split() if t in ranks) for line in u.decode().splitlines()}
                assert st["distinct_baskets"] == len(distinct) and len(distinct) < 5_000
            else:
                assert st["distinct_baskets"] == -1
            recs[dedup] = [r for _, r in got["recommends"]]
        finally:
            sc.close()
    assert recs["1"] == recs["0"]


def test_recommend_second_pass_when_the_trie_walk_stack_overflows(monkeypatch, golden_dir):
    """K9 keeps 160 pending trie ranges per warp in shared memory plus a spill area (4096 by default).  A basket that needs
    more is matched again with a stack sized for the worst case; with the first pass cut down to 24 ranges and no spill area
    (FA_MATCH_SCAP=24, FA_MATCH_GCAP=0) the dense 12-level fixture sends its big baskets down that path, and every
    recommendation must stay what the oracle says."""
    import os
    from oracle import oracle_api
    d = open(os.path.join(golden_dir, "quest_t12_deep", "D.dat"), "rb").read()
    sup = float(open(os.path.join(golden_dir, "quest_t12_deep", "min_support.txt")).read())
    items = sorted({t for line in d.decode().splitlines() for t in line.split()}, key=int)
    rng = np.random.default_rng(2)
    users = [" ".join(items)] * 3 + [" ".join(rng.choice(items, size=len(items) * 3 // 4, replace=False).tolist()) for _ in range(40)]
    users += open(os.path.join(golden_dir, "quest_t12_deep", "U.dat")).read().splitlines()
    u = ("\n".join(users) + "\n").encode()
    ora = oracle_api.run(d, u, sup, fast_rules=True)
    recs = {}
    for gcap in ("0", None):
        sc = _ctx(monkeypatch, FA_MATCH_GCAP=gcap, FA_MATCH_SCAP="24" if gcap else None, FA_USER_DEDUP="0", FA_WEIGHTED=None, FA_K4=None)
        try:
            from helpers import run_product
            got = run_product(sc, d, u, sup)
            st = sc.engine.stats()
            recs[gcap] = [r for _, r in got["recommends"]]
            assert recs[gcap] == ora.recommends
            if gcap == "0":
                assert st["match_retries"] >= 3, "the full baskets must have taken the second pass"
            else:
                assert st["match_retries"] == 0
        finally:
            sc.close()
    assert recs["0"] == recs[None]
