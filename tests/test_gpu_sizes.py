"""Parity at the sizes and shapes BASELINE.json names (tests/test_gpu_parity.py covers the small cases stage by stage).

* configs[0] at full size (100 k rows, min-support 0.092): every level's sets + supports, rule counts, the rule digest,
  the cut counters and the recommendations against the oracle run in this process.
* configs[1] at full size (1 M rows, 0.5 %: 3 M itemsets, 9.8 M rules, 100 k users), configs[3] and configs[4] shapes:
  the result digest against tests/golden/digests.json, which tests/golden/make_digests.py froze from the oracle for the
  same seeds (the oracle needs minutes to hours there, so it ran once, on the CPU box).  bench.py prints the same
  digest for its timed workloads, which ties the timed runs to these checked results.
* configs[2]: ALL pair supports at 16384 items from the tcgen05 kernel against an int32 matrix product."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _golden():
    with open(os.path.join(ROOT, "tests", "golden", "digests.json")) as f:
        return json.load(f)


def _mine(sc, key, rows, users):
    from fastapriori_b200 import AssociationRules, FastApriori, TokenizedLines
    from fastapriori_b200 import utils as Utils
    from fastapriori_b200.workloads import WORKLOADS, gen_rows, gen_users
    wl = WORKLOADS[key]
    off, items = gen_rows(wl, rows, 0, rows)
    uoff, uitems = gen_users(wl, users, 0, users)
    data = TokenizedLines.from_int_csr(off, items, wl["n_items"])
    usr = TokenizedLines.from_int_csr(uoff, uitems, wl["n_items"])
    fa = FastApriori(wl["min_support"], 0)
    fis, i2r, fi = fa.run(sc, data)
    ar = AssociationRules(fis, fi, i2r)
    rec = ar.run(sc, usr, as_arrays=True)
    dig = Utils.runDigest(fis.levels, fi, rec, sc.engine.rules_digest())
    return dict(wl=wl, off=off, items=items, uoff=uoff, uitems=uitems, fa=fa, ar=ar, fis=fis, fi=fi, rec=rec, digest=dig)


def _check_digest(got, want, label):
    n_sets = sum(len(c) for _, c in got["fis"].levels.values())
    detail = (f"{label}: itemsets {n_sets} vs {want['itemsets_n']}, max_k {got['fis'].max_k} vs {want['max_k']}, "
              f"rules {got['ar'].counters.get('rules')} vs {want['rules_n']}, raw {got['ar'].counters.get('raw')} vs {want['raw_rules']}")
    assert got["fa"].counters["minCount"] == want["min_count"], detail
    assert n_sets == want["itemsets_n"] and got["fis"].max_k == want["max_k"], detail
    assert got["digest"]["itemsets"] == want["itemsets"], detail + " (sets or supports differ)"
    assert got["ar"].counters.get("raw") == want["raw_rules"] and got["ar"].counters.get("rules") == want["rules_n"], detail
    assert got["digest"]["rules"] == want["rules"], detail + " (rule multiset or sorted order differs)"
    assert got["digest"]["recommends"] == want["recommends"], detail + " (recommendations differ)"
    assert got["digest"]["all"] == want["all"]


def test_c1_full_size_every_stage_vs_oracle(sc):
    from oracle import oracle_api
    got = _mine(sc, "c1", 100_000, 10_000)
    wl = got["wl"]
    ora = oracle_api.run_csr(got["off"], got["items"], got["uoff"], got["uitems"], wl["min_support"], packed=True, light=False)
    fa, ar, fis = got["fa"], got["ar"], got["fis"]
    assert (fa.counters["count"], fa.counters["minCount"]) == (ora.n_lines, ora.min_count)
    assert got["fi"] == ora.freq_items and len(got["fi"]) >= 20
    assert fis.max_k >= 3
    for k in range(1, max(fis.max_k, max(ora.levels)) + 1):
        o_sets, o_counts = ora.levels.get(k, (np.zeros((0, k), np.int32), np.zeros(0, np.int32)))
        g_sets, g_counts = fis.levels.get(k, (np.zeros((0, k), np.int32), np.zeros(0, np.int32)))
        assert np.array_equal(g_sets, o_sets) and np.array_equal(g_counts, o_counts), f"level {k}"
        if k >= 3 and k in fa.counters:
            assert (fa.counters[k]["groups"], fa.counters[k]["candidates"]) == (ora.n_groups[k], ora.n_cands[k])
    assert (ar.counters["raw"], ar.counters["rules"]) == (ora.n_raw, ora.n_kept)
    for lvl, (b, a) in ar.counters["cut"].items():
        assert (b, a) == (ora.before_cut.get(lvl, 0), ora.after_cut.get(lvl, 0)), f"cut level {lvl}"
    g_ant, g_cons, g_conf = sc.engine.get_rules(1, ar.counters["rules"])
    assert np.array_equal(g_conf.view(np.uint64), ora.rules[2].view(np.uint64)) and np.array_equal(g_cons, ora.rules[1])
    assert sc.engine.rules_digest() == ora.rules_digest
    from fastapriori_b200 import utils as Utils
    assert Utils.recommendsText(got["rec"], got["fi"]) == ora.recommends_text
    assert got["digest"] == Utils.runDigest(ora.levels, ora.freq_items, ora.recommends_text, ora.rules_digest)
    _check_digest(got, _golden()["c1:100000:10000"], "c1")


@pytest.mark.parametrize("name,key,rows,users", [("c2 full size", "c2", 1_000_000, 100_000),
                                                 ("cfg-4 shape", "c4", 4_000_000, 20_000),
                                                 ("cfg-5 shape", "c5", 2_000_000, 40_000)])
def test_digest_matches_oracle_golden(sc, name, key, rows, users):
    gkey = f"{key}:{rows}:{users}"
    golden = _golden()
    if gkey not in golden:
        pytest.skip(f"no oracle digest frozen for {gkey} (tests/golden/make_digests.py)")
    got = _mine(sc, key, rows, users)
    _check_digest(got, golden[gkey], name)
    # independent of the oracle: recount a few supports per level straight from the CSR
    from fastapriori_b200.workloads import WORKLOADS
    n_items = WORKLOADS[key]["n_items"]
    rank_of = np.full(n_items + 1, -1, dtype=np.int64)
    rank_of[np.array([int(s) for s in got["fi"]])] = np.arange(len(got["fi"]))
    ranks = rank_of[got["items"]]
    row_of = np.repeat(np.arange(rows), np.diff(got["off"]))
    rng = np.random.default_rng(1)
    for k, (sets, counts) in got["fis"].levels.items():
        if k < 2 or len(counts) == 0:
            continue
        for i in rng.choice(len(counts), size=min(2, len(counts)), replace=False).tolist():
            hit = np.ones(rows, dtype=bool)
            for x in sets[i].tolist():
                m = np.zeros(rows, dtype=bool)
                m[row_of[ranks == x]] = True
                hit &= m
            assert int(hit.sum()) == int(counts[i]), f"{name}: level {k} row {i}"


def test_pair_matrix_at_16384_items_vs_matrix_product(sc):
    """configs[2] shape: every one of the 134 M pair supports from k_pair_tc (tcgen05 int8, A in TMEM) against
    B^T B computed by an int32 matrix product of the dense 0/1 matrix, and level 2 against the thresholded product."""
    import torch
    eng = sc.engine
    n_items, rows, per_row = 16384, 131_072, 327
    g = torch.Generator(device="cuda").manual_seed(11)
    ids = torch.randint(0, n_items, (rows * per_row,), dtype=torch.int32, device="cuda", generator=g)
    off = torch.arange(0, rows + 1, dtype=torch.int64, device="cuda") * per_row
    eng.attach_transactions(off.data_ptr(), ids.data_ptr(), rows, n_items)
    counts = eng.count_items()
    order = np.argsort(-counts.astype(np.int64), kind="stable").astype(np.int32)
    id_to_rank = np.empty(n_items, dtype=np.int32)
    id_to_rank[order] = np.arange(n_items, dtype=np.int32)
    eng.build_bitmaps(id_to_rank, counts[order])
    min_count = 70
    n2 = eng.count_pairs(min_count)
    ptr, F = eng.pair_counts_device()
    assert F == n_items

    class _Dev:
        def __init__(self, p, n):
            self.__cuda_array_interface__ = {"shape": (n, n), "typestr": "<i4", "data": (p, False), "version": 3, "strides": None}

    P = torch.as_tensor(_Dev(ptr, F), device="cuda")
    # dense 0/1 matrix in RANK space, row chunks of 16384, exact int32 accumulation
    rk = torch.from_numpy(id_to_rank).cuda()[ids.long()].view(rows, per_row).long()
    ref = torch.zeros((F, F), dtype=torch.int32, device="cuda")
    for r0 in range(0, rows, 16384):
        B = torch.zeros((16384, F), dtype=torch.int8, device="cuda")
        B.scatter_(1, rk[r0:r0 + 16384], 1)
        ref += torch._int_mm(B.t().contiguous(), B)
    level1 = torch.diagonal(ref).cpu().numpy()           # presence counts (duplicates inside a row collapse)
    iu = torch.triu(torch.ones((F, F), dtype=torch.bool, device="cuda"), diagonal=1)
    assert torch.equal(P[iu], ref[iu]), "pair supports differ from B^T B"
    assert int((ref[iu] >= min_count).sum()) == n2
    sets2, counts2 = eng.get_level(2)
    want = torch.nonzero((ref >= min_count) & iu).cpu().numpy().astype(np.int32)      # row-major = lexicographic
    assert np.array_equal(sets2, want)
    assert np.array_equal(counts2, ref[want[:, 0], want[:, 1]].cpu().numpy())
    assert n2 > 1000 and level1.max() > 0
