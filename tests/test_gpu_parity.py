"""GPU parity tests: the product path, called through the C-ABI (ctypes -> libfastapriori_b200.so),
against the CPU oracle on the same bytes — bit-exact at every stage."""
import os

import numpy as np
import pytest

from fastapriori_b200 import AssociationRules, FastApriori, NoRulesError, TokenizedLines, quest
from fastapriori_b200 import utils as Utils
from oracle import oracle_api

from helpers import check_against_oracle, run_product

pytestmark = pytest.mark.gpu


def _read(golden_dir, name):
    with open(os.path.join(golden_dir, "appendix_b", name), "rb") as f:
        return f.read()


def _rows_text(rows):
    return ("\n".join(" ".join(map(str, r)) for r in rows) + "\n").encode()


def test_appendix_b_golden(sc, golden_dir):
    d, u = _read(golden_dir, "D.dat"), _read(golden_dir, "U.dat")
    got = run_product(sc, d, u, 0.092)
    assert got["freqitemset_text"] == _read(golden_dir, "freqItemset.part-00000")
    assert b"".join(r.encode() + b"\n" for _, r in got["recommends"]) == _read(golden_dir, "recommends.part-00000")
    n = got["ar"].counters["rules"]
    ant, cons, conf = sc.engine.get_rules(1, n)
    want = [l.split("|") for l in _read(golden_dir, "rules.txt").decode().splitlines() if not l.startswith("#")]
    fi = got["freq_items"]
    have = [[sorted(int(fi[x]) for x in a if x >= 0), int(fi[c]), float(f).hex()] for a, c, f in zip(ant.tolist(), cons.tolist(), conf.tolist())]
    assert have == [[sorted(map(int, a.split())), int(c), float.fromhex(f).hex()] for a, c, f in want]
    golden_rules = b"".join(l + b"\n" for l in _read(golden_dir, "rules.txt").splitlines() if not l.startswith(b"#"))
    assert Utils.rulesLines(sc, got["ar"]) == golden_rules
    check_against_oracle(sc, d, u, 0.092, fast_rules=False)


@pytest.mark.parametrize("seed", range(6))
def test_random_small(sc, seed):
    rng = np.random.default_rng(seed)
    n_items = int(rng.integers(4, 14))
    rows = [rng.integers(1, n_items + 1, size=rng.integers(0, 8)).tolist() for _ in range(int(rng.integers(1, 120)))]
    users = [rng.integers(1, n_items + 3, size=rng.integers(0, 6)).tolist() for _ in range(int(rng.integers(1, 40)))]
    check_against_oracle(sc, _rows_text(rows), _rows_text(users), float(rng.choice([0.03, 0.08, 0.15, 0.3])), fast_rules=False)


def test_edge_tokenisation_duplicates_blank_lines(sc):
    d = b"  1 2\t3 \r\n1  2 2\r\x01 1 3 \n\n1 2 3\n3 1\n\n"
    u = b"1\r\n\n 2 2 \n9 9\n3 1\n"
    check_against_oracle(sc, d, u, 0.3, fast_rules=False)


def test_no_frequent_pairs_is_defined(sc):
    # the reference throws empty.min (AssociationRules.scala:147); declared behaviour here: "0" for all,
    # or NoRulesError in strict mode (what Main uses)
    d, u = b"1\n2\n1\n2\n", b"1\n2 1\n"
    got, ora = check_against_oracle(sc, d, u, 0.5)
    assert ora.status == 1 and [r for _, r in got["recommends"]] == ["0", "0"]
    data = TokenizedLines.from_text(d)
    fis, i2r, fi = FastApriori(0.5).run(sc, data)
    with pytest.raises(NoRulesError):
        AssociationRules(fis, fi, i2r, strict=True).run(sc, TokenizedLines.from_text(u))


def test_nothing_frequent(sc):
    got, ora = check_against_oracle(sc, b"1\n2\n3\n4\n", b"1\n", 0.9)
    assert got["freq_items"] == [] and got["freqitemset_text"] == b""


def test_single_row_and_single_item(sc):
    check_against_oracle(sc, b"5 6 7\n", b"5\n6 7\n", 0.5)
    check_against_oracle(sc, b"5\n5\n5\n", b"5\n", 0.5)


@pytest.mark.parametrize("n_rows,n_items,avg_len,avg_pat,n_pat,sup", [
    (1000, 40, 6, 3, 12, 0.05),      # rows not a multiple of 32 after the first tile
    (4097, 100, 8, 4, 30, 0.02),     # > 4096 rows: second 4 KB chunk of every bitmap; > 64 items: several pair tiles
    (33000, 300, 10, 4, 60, 0.01),   # several bitmap tiles, F > 256
])
def test_quest_mid_size(sc, n_rows, n_items, avg_len, avg_pat, n_pat, sup):
    off, items = quest.generate(n_rows, avg_len, avg_pat, n_pat, n_items, seed=11)
    uoff, uitems = quest.truncate_rows(*quest.generate(300, avg_len, avg_pat, n_pat, n_items, seed=12), seed=5)
    check_against_oracle(sc, quest.to_text(off, items), quest.to_text(uoff, uitems), sup)


def test_c1_style_reference_default_support(sc):
    # BASELINE config 1 shape at reduced rows: T10I4-style, 1000 items, the reference's hard-coded 0.092
    off, items = quest.generate(20000, 10, 4, 30, 1000, seed=1)
    uoff, uitems = quest.truncate_rows(*quest.generate(500, 10, 4, 30, 1000, seed=101), seed=1)
    got, ora = check_against_oracle(sc, quest.to_text(off, items), quest.to_text(uoff, uitems), 0.092)
    assert len(ora.levels[2][1]) > 0, "generator must yield at least one frequent pair at 9.2% (A-16)"


def test_deep_levels(sc):
    # one heavy 12-item pattern: 2^12 frequent subsets, 12 levels, tens of thousands of rules
    rng = np.random.default_rng(3)
    pat = list(range(1, 13))
    rows = []
    for _ in range(600):
        r = [x for x in pat if rng.random() < 0.93] + rng.integers(13, 40, size=3).tolist()
        rows.append(r)
    users = [[x for x in pat if rng.random() < 0.5] + [int(rng.integers(13, 40))] for _ in range(200)]
    got, ora = check_against_oracle(sc, _rows_text(rows), _rows_text(users), 0.3)
    assert got["fis"].max_k >= 10


def test_row_trimming_keeps_supports_exact(sc):
    # a heavy 10-item pattern lives in ~8 % of 30k rows: after a few levels only those rows can support a
    # candidate, so the level loop squeezes the bitmaps; every support must still match the oracle
    rng = np.random.default_rng(8)
    pat = list(range(1, 11))
    rows = []
    for _ in range(30000):
        if rng.random() < 0.08:
            rows.append([x for x in pat if rng.random() < 0.95] + rng.integers(11, 60, size=4).tolist())
        else:
            rows.append(rng.integers(11, 60, size=6).tolist() + [int(rng.integers(1, 11))])
    users = [[x for x in pat if rng.random() < 0.6] + [int(rng.integers(11, 60))] for _ in range(300)]
    sc.engine.reset_stats()
    got, ora = check_against_oracle(sc, _rows_text(rows), _rows_text(users), 0.03)
    st = sc.engine.stats()
    assert got["fis"].max_k >= 8
    assert st["row_trims"] >= 1 and st["live_rows"] < 30000 * 0.7


@pytest.mark.parametrize("n_items", [2000, 4000, 7000])
def test_many_items_rank_chunks(sc, n_items):
    # Large F: 7000 forces K2 to split the ranks over several shared-memory passes; 2000 / 4000 / 7000 make the
    # column form of K4 run with 4, 2 and 1 lanes per unit.  The oracle's pair pass is O(F^2 * rows) on the CPU,
    # so these cases are checked against direct recounts from the rows instead.
    rng = np.random.default_rng(5)
    n_rows = 3000
    rows = [np.concatenate(([1, 2, 3], 4 + rng.choice(n_items - 3, size=12, replace=False))).tolist() for _ in range(n_rows)]
    data = TokenizedLines.from_lines([[str(x) for x in r] for r in rows])
    fa = FastApriori(1.5 / n_rows)
    fis, i2r, fi = fa.run(sc, data)
    assert fa.counters["minCount"] == 2
    F = len(fi)
    assert F > 0.9 * n_items
    flat = np.array([x for r in rows for x in r])
    occ = np.bincount(flat, minlength=n_items + 1)
    assert F == int((occ >= 2).sum())
    assert fis.levels[1][1].tolist() == sorted(occ[occ >= 2].tolist(), reverse=True)
    # items 1,2,3 sit in every row: support({1, x}) == count(x) for every frequent x
    sets2, counts2 = fis.levels[2]
    r1 = i2r["1"]
    c1 = fis.levels[1][1]
    with_1 = (sets2[:, 0] == r1) | (sets2[:, 1] == r1)
    other = np.where(sets2[:, 0] == r1, sets2[:, 1], sets2[:, 0])[with_1]
    assert with_1.sum() == F - 1
    assert np.array_equal(counts2[with_1], c1[other])
    # every pair is recounted from the rows
    pair_count = {}
    rank_rows = [sorted(i2r[str(x)] for x in r if str(x) in i2r) for r in rows]
    for r in rank_rows:
        for i in range(len(r)):
            for j in range(i + 1, len(r)):
                pair_count[(r[i], r[j])] = pair_count.get((r[i], r[j]), 0) + 1
    want = {p: c for p, c in pair_count.items() if c >= 2}
    assert {tuple(p): c for p, c in zip(sets2.tolist(), counts2.tolist())} == want
    # level 3 recounted from the rows as well (exercises K4 with this F)
    tri = {}
    for r in rank_rows:
        n = len(r)
        for i in range(n):
            for j in range(i + 1, n):
                for l in range(j + 1, n):
                    t = (r[i], r[j], r[l])
                    tri[t] = tri.get(t, 0) + 1
    sets3, counts3 = fis.levels[3]
    assert {tuple(p): c for p, c in zip(sets3.tolist(), counts3.tolist())} == {t: c for t, c in tri.items() if c >= 2}
    sets4, counts4 = fis.levels[4]
    assert len(counts4) >= F - 3 and np.all(counts4 >= 2)


@pytest.mark.parametrize("n_rows,n_items", [(20000, 300), (5000, 129), (70000, 520)])
def test_pair_pass_tensor_core_equals_popc(sc, monkeypatch, n_rows, n_items):
    # K3 has two kernels: tcgen05 int8 GEMM (default from 256 frequent items) and the AND+POPC kernel.  Both must give
    # the same frequent pairs and supports; the POPC one is checked against the oracle by the tests above.
    off, items = quest.generate(n_rows, 12, 5, 60, n_items, seed=17)
    data = TokenizedLines.from_int_csr(off, items, n_items)
    out = {}
    impls = ("popc", "tc", "tc256", "tc512", "tcp256")
    # "tc": 128 x 128 items per CTA (k_pair_tc); "tc256": 256 x 128 (k_pair_tc2); "tc512" / "tcp256": CTA pairs with
    # cta_group::2 MMAs, two N = 128 accumulators (k_pair_tc3) / one N = 256 accumulator (k_pair_tc4, the default above 1024 items)
    for impl in impls:
        monkeypatch.setenv("FA_PAIR_IMPL", impl[:2] if impl != "popc" else impl)
        monkeypatch.setenv("FA_PAIR_TILE", impl[2:] if len(impl) > 2 else "128")
        fis, _, fi = FastApriori(0.002).run(sc, data)
        out[impl] = (fi, fis.levels[2][0].copy(), fis.levels[2][1].copy(), fis.as_dict())
    for pair_impl in ("tc512", "tcp256"):   # the CTA-pair kernels (cta_group::2)
        if pair_impl in out:
            assert np.array_equal(out[pair_impl][1], out["popc"][1]) and np.array_equal(out[pair_impl][2], out["popc"][2]), pair_impl
    assert out["tc256"][0] == out["popc"][0] and out["tc256"][3] == out["popc"][3]
    assert np.array_equal(out["tc256"][1], out["popc"][1]) and np.array_equal(out["tc256"][2], out["popc"][2])
    assert out["tc"][0] == out["popc"][0]
    assert len(out["popc"][2]) > 100
    assert np.array_equal(out["tc"][1], out["popc"][1]) and np.array_equal(out["tc"][2], out["popc"][2])
    assert out["tc"][3] == out["popc"][3]
    # and against a direct recount of a sample of pairs from the rows
    ranks = {s: r for r, s in enumerate(out["tc"][0])}
    rows = [set(ranks[str(x)] for x in items[off[r]:off[r + 1]] if str(x) in ranks) for r in range(n_rows)]
    rng = np.random.default_rng(0)
    for idx in rng.choice(len(out["tc"][2]), size=25, replace=False).tolist():
        a, b = out["tc"][1][idx].tolist()
        assert sum(1 for rs in rows if a in rs and b in rs) == int(out["tc"][2][idx])


def test_resident_csr_equals_host_csr(sc):
    import torch
    off, items = quest.generate(5000, 8, 4, 20, 60, seed=21)
    data = TokenizedLines.from_int_csr(off, items, 60)
    a, _, fa_items = FastApriori(0.03).run(sc, data)
    d_off = torch.from_numpy(data.offsets).cuda()
    d_ids = torch.from_numpy(data.ids).cuda()
    b, _, fb_items = FastApriori(0.03).run(sc, data, resident=(d_off.data_ptr(), d_ids.data_ptr()))
    assert fa_items == fb_items and a.as_dict() == b.as_dict()


def test_rule_order_with_confidences_one_gap_apart(sc):
    # K8 sorts the confidence keys on the bit range that can differ, derived from the support bounds (fa_sort_rules):
    # confidences (n-1)/n, (n-2)/(n-1), (n-3)/(n-2) ... are the closest two rules can get for supports <= n
    n = 4000
    rows = []
    for r in range(n):
        row = [1]
        if r != 0: row.append(2)
        if r != 1: row.append(3)
        if r not in (1, 2): row.append(4)
        if r not in (1, 2, 3): row.append(5)
        if r % 2: row.append(6)
        rows.append(row)
    users = [[1], [3], [4], [2, 6], [5], [1, 2, 3, 4], [6]]
    for ms in (0.4, 0.9):
        check_against_oracle(sc, _rows_text(rows), _rows_text(users), ms, fast_rules=False)


def test_gen_candidates_waited_form_equals_queued_form(sc, monkeypatch):
    # fa_gen_candidates with outputs waits for K5's totals and sizes K4 exactly; with two null pointers (what the host
    # loop passes) everything is sized by host-side bounds and the totals come back with K4's: same levels, same counters
    off, items = quest.generate(6000, 9, 4, 25, 70, seed=33)
    data = TokenizedLines.from_int_csr(off, items, 70)
    fa1 = FastApriori(0.02)
    a, _, _ = fa1.run(sc, data)
    eng = sc.engine
    queued = eng.gen_candidates
    seen = {}

    def waited(k, wait=True):
        seen[k] = queued(k, True)
        return None

    monkeypatch.setattr(eng, "gen_candidates", waited)
    fa2 = FastApriori(0.02)
    b, _, _ = fa2.run(sc, data)
    assert a.as_dict() == b.as_dict()
    assert fa1.counters == fa2.counters and fa1.counters["max_k"] >= 4
    for k, (g, c) in seen.items():
        assert (g, c) == (fa1.counters[k]["groups"], fa1.counters[k]["candidates"]) == eng.level_counters(k)
    assert eng.level_counters(2) == (0, 0) and eng.level_counters(60) == (0, 0)
    from fastapriori_b200._native import NativeError
    import ctypes
    with pytest.raises(NativeError):   # one output only
        eng._check(eng._lib.fa_gen_candidates(eng._ctx, 3, ctypes.byref(ctypes.c_int64()), None))


def test_compact_16bit_ids_give_the_same_lattice(sc):
    # fa_load_transactions_u16: the ids cross PCIe as uint16 and are widened on the device
    off, items = quest.generate(7001, 8, 4, 20, 60, seed=23)   # odd element count: exercises the widening kernel's tail
    data = TokenizedLines.from_int_csr(off, items, 60)
    small = data.compact()
    assert small.ids.dtype == np.uint16 and np.array_equal(small.ids, data.ids)
    a, _, fa_items = FastApriori(0.03).run(sc, data)
    b, _, fb_items = FastApriori(0.03).run(sc, small)
    assert fa_items == fb_items and a.as_dict() == b.as_dict() and len(a.as_dict()) > 60
    from fastapriori_b200._native import NativeError
    with pytest.raises(NativeError):
        sc.engine.load_transactions(np.array([0, 1], dtype=np.int64), np.array([3], dtype=np.uint16), 70000)


def test_bad_inputs_are_rejected(sc):
    from fastapriori_b200._native import NativeError
    eng = sc.engine
    with pytest.raises(NativeError):
        eng.load_transactions([0, 2], [0, 7], 3)  # id 7 outside [0, 3)
        eng.count_items()
    with pytest.raises(NativeError):
        eng.gen_candidates(7)                      # levels 2..6 do not exist
    eng.load_transactions([0, 2], [0, 1], 2)
    with pytest.raises(NativeError):
        eng.count_pairs(1)                         # bitmaps not built yet


def test_main_contract(sc, tmp_path, golden_dir):
    # Main.scala:24-37: <input>/<output> are prefixes; outputs are directories with part-00000 + _SUCCESS
    from fastapriori_b200 import main as Main
    inp = str(tmp_path) + "/in_"
    out = str(tmp_path) + "/out_"
    for n in ("D.dat", "U.dat"):
        with open(inp + n, "wb") as f:
            f.write(_read(golden_dir, n))
    Main.main([inp, out, str(tmp_path / "tmp")])
    assert open(out + "freqItemset/part-00000", "rb").read() == _read(golden_dir, "freqItemset.part-00000")
    assert open(out + "recommends/part-00000", "rb").read() == _read(golden_dir, "recommends.part-00000")
    assert os.path.exists(out + "freqItemset/_SUCCESS") and os.path.exists(out + "recommends/_SUCCESS")
    with pytest.raises(FileExistsError):          # saveAsTextFile refuses to overwrite (Utils.scala:39)
        Main.main([inp, out, "ignored"])


def test_cpp_host_twin(sc, tmp_path, golden_dir):
    # fa_main: the same C-ABI sequence from C++ (no Python in that process); outputs must be byte-identical
    import subprocess
    from fastapriori_b200 import build
    exe = build.build_main()
    inp, out = str(tmp_path) + "/in_", str(tmp_path) + "/cpp_"
    for n in ("D.dat", "U.dat"):
        with open(inp + n, "wb") as f:
            f.write(_read(golden_dir, n))
    r = subprocess.run([exe, inp, out, "tmp"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "==== 3 candidate items 1" in r.stdout and "==== Size association rules 7" in r.stdout
    assert open(out + "freqItemset/part-00000", "rb").read() == _read(golden_dir, "freqItemset.part-00000")
    assert open(out + "recommends/part-00000", "rb").read() == _read(golden_dir, "recommends.part-00000")
    # a larger case against the oracle, with a non-default min support
    off, items = quest.generate(8000, 10, 4, 40, 150, seed=41)
    uoff, uitems = quest.truncate_rows(*quest.generate(400, 10, 4, 40, 150, seed=42), seed=6)
    d, u = quest.to_text(off, items), quest.to_text(uoff, uitems)
    inp2, out2 = str(tmp_path) + "/in2_", str(tmp_path) + "/cpp2_"
    open(inp2 + "D.dat", "wb").write(d)
    open(inp2 + "U.dat", "wb").write(u)
    r = subprocess.run([exe, inp2, out2, "tmp", "0.02"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    ora = oracle_api.run(d, u, 0.02, fast_rules=True)
    assert open(out2 + "freqItemset/part-00000", "rb").read() == ora.freqitemset_text
    assert open(out2 + "recommends/part-00000", "rb").read() == ora.recommends_text


@pytest.mark.parametrize("name", ["quest_t8", "quest_t12_deep"])
def test_committed_golden_fixtures(sc, golden_dir, name):
    # product vs bytes committed under tests/golden/ (made by tests/golden/make_quest_golden.py); no oracle involved here
    d = os.path.join(golden_dir, name)
    rd = lambda f: open(os.path.join(d, f), "rb").read()
    got = run_product(sc, rd("D.dat"), rd("U.dat"), float(rd("min_support.txt")))
    assert got["freqitemset_text"] == rd("freqItemset.part-00000")
    assert b"".join(r.encode() + b"\n" for _, r in got["recommends"]) == rd("recommends.part-00000")
    assert Utils.freqItemsetLines(got["fis"], got["freq_items"], with_count=True) == rd("counts.txt")
    mine = Utils.rulesLines(sc, got["ar"]).splitlines()
    want = rd("rules.txt").splitlines()
    key = lambda l: l.split(b"|")[1:]                      # (consequent, confidence): the order the reference defines
    assert [key(l) for l in mine] == [key(l) for l in want]
    assert sorted(mine) == sorted(want)                    # same rules; order inside full ties is free (A-13)


def test_resume_rule_stage_from_counts_dump(sc, tmp_path):
    # items[count] dump -> loader -> fa_put_level -> the rule stage gives what the mining run gave
    from fastapriori_b200 import MiningContext
    from fastapriori_b200._native import NativeError
    off, items = quest.generate(30_000, 10, 5, 60, 120, seed=31)
    data = TokenizedLines.from_int_csr(off, items, 120)
    uoff, uitems = quest.truncate_rows(*quest.generate(3000, 10, 5, 60, 120, seed=32), seed=3)
    users = TokenizedLines.from_int_csr(uoff, uitems, 120)
    fis, i2r, fi = FastApriori(0.01).run(sc, data)
    ar = AssociationRules(fis, fi, i2r)
    rec = ar.run(sc, users, as_arrays=True)
    rules = Utils.rulesLines(sc, ar)
    assert fis.max_k >= 4 and ar.counters["rules"] > 100
    out = str(tmp_path) + "/"
    Utils.saveFreqItemsetWithCount(sc, out, fis, fi)

    fis2, i2r2, fi2 = Utils.loadFreqItemsetWithCount(out + "freqItems")
    assert fi2 == fi and i2r2 == i2r and fis2.as_dict() == fis.as_dict()
    other = MiningContext(device=sc.device, distributed=False)  # a context that never saw the transactions
    try:
        ar2 = AssociationRules(fis2, fi2, i2r2)
        rec2 = ar2.run(other, users, as_arrays=True)
        assert np.array_equal(rec2, rec)
        assert ar2.counters["rules"] == ar.counters["rules"] and ar2.counters["cut"] == ar.counters["cut"]
        assert Utils.rulesLines(other, ar2) == rules
        # a lattice that is not downward closed, or not sorted, is refused (the rule kernels index subsets)
        lv = {k: (s.copy(), c.copy()) for k, (s, c) in fis2.levels.items()}
        lv[2] = (lv[2][0][1:], lv[2][1][1:])
        with pytest.raises(NativeError, match="downward closed"):
            other.engine.put_levels(lv)
        lv = {k: (s.copy(), c.copy()) for k, (s, c) in fis2.levels.items()}
        lv[3] = (lv[3][0][::-1].copy(), lv[3][1][::-1].copy())
        with pytest.raises(NativeError, match="lexicographic"):
            other.engine.put_levels(lv)
    finally:
        other.close()
    # the mining context is untouched: its own itemsets are still resident and usable
    assert np.array_equal(AssociationRules(fis, fi, i2r).run(sc, users, as_arrays=True), rec)
