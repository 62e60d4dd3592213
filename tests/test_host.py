"""CPU tests of the host mirror: tokenisation, rank order, basket remap, output formatting, the C-ABI
surface (library loads and exports every declared symbol) and the no-fallback rule."""
import ctypes
import os
import re

import numpy as np
import pytest

from fastapriori_b200 import TokenizedLines, _native
from fastapriori_b200 import utils as Utils
from fastapriori_b200.association_rules import AssociationRules
from fastapriori_b200.fast_apriori import FreqItemsets, _int_or_none, rank_order
from oracle import oracle_api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tokeniser_matches_oracle_rules():
    d = b"  1 2\t3 \r\n1  2\r\x01 1 \n\n1 2\n"
    t = TokenizedLines.from_text(d)
    assert len(t) == 5
    assert [t.line(i) for i in range(5)] == [["1", "2", "3"], ["1", "2"], ["1"], [""], ["1", "2"]]
    o = oracle_api.run(d, b"", 0.5, skip_rules=True)
    counts = np.bincount(t.ids, minlength=len(t.vocab))
    order = rank_order(t.vocab, counts, o.min_count)
    assert [t.vocab[i] for i in order] == o.freq_items
    assert TokenizedLines.from_text(b"").count() == 0
    assert TokenizedLines.from_text(b"7").count() == 1          # no trailing newline still yields the record
    assert TokenizedLines.from_text(b"7\n\n").count() == 2      # ...and an inner blank line is a record


def test_native_loader_matches_python_rules():
    rng = np.random.default_rng(3)
    cases = [b"", b"7", b"7\n\n", b"  1 2\t3 \r\n1  2\r\x01 1 \n\n1 2\n", b"\n\n\n", b"a\x0bb\x0cc d\r\r\n e",
             b"x\x01y z\n", "caf\u00e9 na\u00efve\n".encode()]
    for _ in range(30):
        toks = [b"1", b"22", b"333", b"x", b" ", b"  ", b"\t", b"\n", b"\r\n", b"\r", b"\x0b", b"\x01"]
        cases.append(b"".join(toks[i] for i in rng.integers(0, len(toks), size=rng.integers(0, 60))))
    for c in cases:
        a, b = TokenizedLines.from_text(c), TokenizedLines.from_text_py(c)
        assert len(a) == len(b), c
        assert [a.line(i) for i in range(len(a))] == [b.line(i) for i in range(len(b))], c
        assert a.vocab == b.vocab and np.array_equal(a.offsets, b.offsets) and np.array_equal(a.ids, b.ids), c


def test_compact_ids():
    d = TokenizedLines.from_lines([["a", "b"], [], ["b", "c", "a"]])
    c = d.compact()
    assert c.ids.dtype == np.uint16 and c.ids.tolist() == d.ids.tolist() and c.vocab == d.vocab
    assert [c.line(i) for i in range(len(c))] == [["a", "b"], [], ["b", "c", "a"]]
    out = np.zeros(len(d.ids), dtype=np.uint16)
    assert d.compact(out=out).ids is out or np.shares_memory(d.compact(out=out).ids, out)
    big = TokenizedLines(np.array([0, 1]), np.array([70000], dtype=np.int32), [str(i) for i in range(70001)])
    assert big.compact() is big


def _conf_key_bits(lo, hi):
    lib = _native.load()
    a, b = ctypes.c_int32(), ctypes.c_int32()
    assert lib.fa_conf_key_bits(int(lo), int(hi), ctypes.byref(a), ctypes.byref(b)) == 0
    return a.value, b.value


def _truncated_keys(conf, lo_bit, hi_bit):
    key = ~conf.view(np.uint64)                                   # K8's primary key: confidence descending
    return (key >> np.uint64(lo_bit)) & np.uint64((1 << (hi_bit - lo_bit)) - 1)


def test_confidence_key_bit_range_orders_every_pair_of_confidences():
    # fa_sort_rules sorts ~bits(conf) on [lo_bit, hi_bit) only.  Exhaustive over all supports a <= b in [sup_min, sup_max] for
    # small bounds, and over the closest possible confidences ((b-1)/b against (b-2)/(b-1), ...) near large bounds:
    # different confidences must get different truncated keys in the same order, equal confidences equal keys.
    assert _conf_key_bits(0, 10) == (0, 64) and _conf_key_bits(5, 4) == (0, 64) and _conf_key_bits(-1, 7) == (0, 64)
    for sup_max in (1, 2, 3, 7, 8, 33, 100, 257, 1000):
        for sup_min in sorted({1, 2, max(1, sup_max // 3), sup_max}):
            if sup_min > sup_max:
                continue
            lo_bit, hi_bit = _conf_key_bits(sup_min, sup_max)
            assert 0 <= lo_bit < hi_bit <= 64
            a, b = np.meshgrid(np.arange(sup_min, sup_max + 1), np.arange(sup_min, sup_max + 1))
            m = a <= b
            conf = np.unique(a[m].astype(np.float64) / b[m].astype(np.float64))      # ascending, distinct
            t = _truncated_keys(conf, lo_bit, hi_bit)
            assert np.all(t[1:] < t[:-1]), (sup_min, sup_max)                        # keys descend as confidences ascend
    rng = np.random.default_rng(5)
    for sup_max in (4093, 65536, 1 << 20, (1 << 24) + 3, (1 << 27) - 1, (1 << 30) + 12345):
        lo_bit, hi_bit = _conf_key_bits(1, sup_max)
        b = np.concatenate([np.arange(sup_max - 400, sup_max + 1), rng.integers(2, sup_max, 4000)]).astype(np.int64)
        num = np.concatenate([b - 1, b - 2, np.maximum(b // 2, 1), np.ones_like(b), b])
        den = np.concatenate([b, b, b, b, b])
        num = np.concatenate([num, np.maximum(num - 1, 1)])
        den = np.concatenate([den, np.maximum(den - 1, 1)])
        ok = (num >= 1) & (num <= den)
        conf = np.unique(num[ok].astype(np.float64) / den[ok].astype(np.float64))
        t = _truncated_keys(conf, lo_bit, hi_bit)
        assert np.all(t[1:] < t[:-1]), sup_max
        if sup_max < (1 << 24):
            assert hi_bit - lo_bit < 64                                               # and the range really is narrower


def test_rank_order_canonical_ties():
    vocab = ["10", "9", "x", "-3", "b", "007"]
    counts = np.array([5, 5, 5, 5, 5, 9])
    order = rank_order(vocab, counts, 5)
    # count desc; ties: ints by value (-3, 9, 10), then non-ints in byte order
    assert [vocab[i] for i in order] == ["007", "-3", "9", "10", "b", "x"]
    assert _int_or_none("+12") == 12 and _int_or_none("2147483648") is None and _int_or_none("1.5") is None
    assert len(rank_order(vocab, counts, 10)) == 0


def test_freq_itemset_lines_format():
    fis = FreqItemsets({1: (np.array([[0], [1], [2]], np.int32), np.array([21, 14, 13], np.int32)),
                        2: (np.array([[0, 1], [0, 2], [1, 2]], np.int32), np.array([12, 12, 6], np.int32)),
                        3: (np.array([[0, 1, 2]], np.int32), np.array([6], np.int32))})
    items = ["10", "30", "20"]
    assert Utils.freqItemsetLines(fis, items) == b"10\n20\n20 10\n20 30\n20 30 10\n30\n30 10\n"      # Utils.scala:38-39
    assert Utils.freqItemsetLines(fis, items, with_count=True).splitlines()[2] == b"20 30 10[6]"     # Utils.scala:60
    assert len(fis) == 7 and list(fis)[0] == (frozenset({0, 1}), 12) and list(fis)[-1] == (frozenset({2}), 13)


def test_remove_redundancy_maps_and_filters():
    fis = FreqItemsets({}, engine=object())
    ar = AssociationRules(fis, ["10", "30", "20"], {"10": 0, "30": 1, "20": 2})
    users = TokenizedLines.from_text(b"20 30\n40 99\n\n10 10 20\n")
    off, ranks = ar.removeRedundancy(None, users)
    assert off.tolist() == users.offsets.tolist() and ranks.tolist() == [2, 1, -1, -1, -1, 0, 0, 2]   # 40, 99 and "" (the blank line's token) map to -1
    with pytest.raises(TypeError):
        AssociationRules([(frozenset({0}), 1)], ["10"], {"10": 0})


def test_header_symbols_are_exported():
    hdr = open(os.path.join(ROOT, "include", "fastapriori_b200.h")).read()
    declared = set(re.findall(r"^(?:int|void|const char\*)\s+(fa_\w+)\s*\(", hdr, flags=re.M))
    assert len(declared) >= 20
    lib = _native.load()            # builds in-tree if needed; raises if a bound symbol is missing
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in the header but not exported"
    assert declared <= set(_native.SIGNATURES), "every declared entry point has a ctypes signature"
    # error behaviour that needs no GPU: null ctx / null out pointers
    assert lib.fa_create(0, None, None) == 1
    assert lib.fa_last_error(None) == b"null ctx"
    assert lib.fa_count_items(None, None) == 1


def test_no_cpu_fallback_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from fastapriori_b200 import MiningContext
    from fastapriori_b200.engine import Engine
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        MiningContext()
    with pytest.raises(_native.NativeError):
        Engine(0)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "fastapriori_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".scala")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in src.replace("the oracle", "").replace("(and by the oracle)", ""), f"{f} mentions oracle/"


def test_quest_generator_is_shard_invariant():
    from fastapriori_b200 import quest
    off, items = quest.generate(40000, 10, 4, 50, 200, seed=3)
    off2, items2 = quest.generate(40000, 10, 4, 50, 200, seed=3, row_begin=17000, row_end=35000)
    a, b = off[17000], off[35000]
    assert np.array_equal(items[a:b], items2) and np.array_equal(off[17000:35001] - a, off2)
    assert items.min() >= 1 and items.max() <= 200
    lens = np.diff(off)
    assert 8 < lens.mean() < 12
    for r in range(0, 40000, 997):                      # items are distinct within a row (A-2)
        row = items[off[r]:off[r + 1]]
        assert len(set(row.tolist())) == len(row)
    uo, ui = quest.truncate_rows(off[:101], items[:off[100]], seed=1)
    assert np.all(np.diff(uo) >= 1) and np.all(np.diff(uo) <= lens[:100])


def test_counts_dump_round_trip():
    # Utils.saveFreqItemsetWithCount format (Utils.scala:51-63) -> loadFreqItemsetWithCount: same lattice, same ranks
    from fastapriori_b200 import FreqItemsets
    fi = ["7", "3", "12", "x y".replace(" ", "_"), "5"]          # rank order: counts 9, 8, 8, 6, 5 (3 before 12: tie rule)
    levels = {1: (np.arange(5, dtype=np.int32).reshape(-1, 1), np.array([9, 8, 8, 6, 5], dtype=np.int32)),
              2: (np.array([[0, 1], [0, 2], [1, 2], [1, 4]], dtype=np.int32), np.array([7, 6, 5, 4], dtype=np.int32)),
              3: (np.array([[0, 1, 2]], dtype=np.int32), np.array([4], dtype=np.int32))}
    text = Utils._freqItemsetLines_slow(FreqItemsets(levels), [s.encode() for s in fi], True)
    assert b"12 3 7[4]\n" in text and b"7[9]\n" in text
    fis, i2r, fi2 = Utils.loadFreqItemsetWithCount(text)
    assert fi2 == fi and i2r == {s: r for r, s in enumerate(fi)}
    assert fis.as_dict() == FreqItemsets(levels).as_dict()
    for k in levels:
        assert np.array_equal(fis.levels[k][0], levels[k][0]) and np.array_equal(fis.levels[k][1], levels[k][1])
    with pytest.raises(ValueError):
        Utils.loadFreqItemsetWithCount(b"1 2 3\n")


def test_spark_tie_order_emulation():
    # SURVEY.md Appendix C / 8 f-2.  Known answers: String.hashCode, Guava murmur3_32.hashInt (cross-checked against
    # scikit-learn's MurmurHash3_x86_32 on the 4 little-endian bytes), then the partition / slot order built on them.
    from fastapriori_b200.fast_apriori import java_string_hash, murmur3_int, spark_collect_order
    assert [java_string_hash(s) for s in ("", "a", "abc", "1000", "polygenelubricants")] == [0, 97, 96354, 1507423, -2147483648]
    assert java_string_hash("\U0001F600") == (0xD83D * 31 + 0xDE00)            # UTF-16 code units, not code points
    assert murmur3_int(0) & 0xFFFFFFFF == 0x2362F9DE
    sk = pytest.importorskip("sklearn.utils")
    for h in (0, 1, -1, 97, 96354, 1507423, -2147483648, 2147483647, 123456789):
        assert murmur3_int(h) == sk.murmurhash3_32(int(h & 0xFFFFFFFF).to_bytes(4, "little"), seed=0, positive=False)
    # one-character digits: hashCode 49..57, one per partition -> collect() order is partition order
    assert spark_collect_order(list("987654321"), 500).tolist() == [8, 7, 6, 5, 4, 3, 2, 1, 0]
    # a single partition: slot order of one AppendOnlyMap, independent of insertion order while nothing collides
    toks = [str(i) for i in range(1, 40)]
    one = spark_collect_order(toks, 1)
    slots = {t: murmur3_int(java_string_hash(t)) & 63 for t in toks}
    if len(set(slots.values())) == len(toks):
        assert [toks[i] for i in np.argsort(one)] == sorted(toks, key=slots.get)
    assert sorted(one.tolist()) == list(range(len(toks)))
    many = spark_collect_order([str(i) for i in range(1, 301)], 1)               # 300 keys: the table grows 64 -> 512
    assert sorted(many.tolist()) == list(range(300))
    # rank order: count descending first, the emulated order only among equal counts
    vocab = ["5", "17", "3", "250"]
    counts = np.array([4, 9, 4, 4], dtype=np.int32)
    got = rank_order(vocab, counts, 1, "spark", 500).tolist()
    assert got[0] == 1 and sorted(got[1:]) == [0, 2, 3]
    parts = {i: java_string_hash(vocab[i]) % 500 for i in (0, 2, 3)}
    assert got[1:] == sorted((0, 2, 3), key=parts.get)                           # 3 distinct partitions here
    assert rank_order(vocab, counts, 1).tolist() == [1, 2, 0, 3]                 # canonical: Int value ascending


def test_native_writer_matches_python_sort_randomised():
    # fa_lines_* (keyed bucket sort, parallel) against the explicit Python formatter + UTF-16 sort, incl. short lines,
    # shared 8-byte prefixes and the [count] suffix
    rng = np.random.default_rng(5)
    for trial in range(8):
        F = int(rng.integers(1, 40))
        items = [str(int(x) * int(rng.choice([1, 7, 1234567]))) for x in rng.permutation(np.arange(1, F + 1))]
        if trial % 3 == 0:
            items = ["a" * int(rng.integers(1, 12)) + str(i) for i in range(F)]
        levels = {1: (np.arange(F, dtype=np.int32).reshape(-1, 1), rng.integers(1, 100, F).astype(np.int32))}
        for k in range(2, min(F, 6) + 1):
            n = int(rng.integers(0, 1500))
            rows = np.array([np.sort(rng.choice(F, size=k, replace=False)) for _ in range(n)], dtype=np.int32).reshape(-1, k)
            levels[k] = (rows, rng.integers(1, 100, n).astype(np.int32))
        fis = FreqItemsets(levels)
        for with_count in (False, True):
            assert Utils.freqItemsetLines(fis, items, with_count=with_count) == \
                Utils._freqItemsetLines_slow(fis, [s.encode() for s in items], with_count)


def test_native_loader_multi_chunk_growing_vocabulary():
    # > 1 MB per chunk, so the loader really splits the input; the vocabulary keeps growing across chunks and includes
    # tokens of exactly 8, 9 and more bytes (packed-key table vs the string map), numbering must stay first-seen
    lines = []
    for i in range(150_000):
        toks = [str((i * 7 + j) % 30_000) for j in range(10)]
        if i % 97 == 0:
            toks += ["12345678", "123456789", "x" * (9 + i % 5)]
        lines.append(" ".join(toks))
    buf = ("\n".join(lines) + "\n").encode()
    assert len(buf) > 8 << 20
    a, b = TokenizedLines.from_text(buf), TokenizedLines.from_text_py(buf)
    assert a.vocab == b.vocab and np.array_equal(a.offsets, b.offsets) and np.array_equal(a.ids, b.ids)


def test_spark_tie_order_through_the_operator():
    # FastApriori(tie_order="spark") on the host logic alone (numpy stand-in engine): the same itemsets and supports as
    # the canonical rule, only the ranks of equal-count items - hence the item order inside a saved line - may differ
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from fake_engine import FakeEngine
    from fastapriori_b200 import FastApriori, MiningContext
    from fastapriori_b200.fast_apriori import spark_collect_order
    text = b"5 17 3\n17 3 250\n5 250 17\n3 5 250\n17 5 3 250\n"          # every item occurs 4 times: one big tie
    data = TokenizedLines.from_text(text)
    runs = {}
    for mode in ("canonical", "spark"):
        eng = FakeEngine()
        sc = MiningContext(engine=eng, distributed=False)
        eng.allreduce = sc.allreduce_host
        fis, i2r, fi = FastApriori(0.4, 500, tie_order=mode).run(sc, data)
        named = {frozenset(fi[r] for r in row): c for row, c in fis.as_dict().items()}
        runs[mode] = (fi, named, Utils.freqItemsetLines(fis, fi))
    assert runs["canonical"][0] == ["3", "5", "17", "250"]
    pos = spark_collect_order(data.vocab, 500)
    assert runs["spark"][0] == [data.vocab[i] for i in np.argsort(pos)]
    assert runs["spark"][1] == runs["canonical"][1] and len(runs["spark"][1]) > 8
    assert sorted(map(sorted, (l.split() for l in runs["spark"][2].splitlines()))) == \
        sorted(map(sorted, (l.split() for l in runs["canonical"][2].splitlines())))


def test_workload_users_are_shard_invariant():
    # bench.py shards the users of a workload over the ranks: basket u must not depend on how the range was cut
    from fastapriori_b200.workloads import WORKLOADS, gen_rows, gen_users
    wl = WORKLOADS["c5"]
    off, items = gen_users(wl, 3000, 0, 3000)
    a = gen_users(wl, 3000, 0, 1234)
    b = gen_users(wl, 3000, 1234, 3000)
    assert np.array_equal(items, np.concatenate([a[1], b[1]]))
    assert np.array_equal(np.diff(off), np.concatenate([np.diff(a[0]), np.diff(b[0])]))
    assert np.diff(off).min() >= 1 and items.min() >= 1 and items.max() <= wl["n_items"]
    # and the rows of a workload do not depend on the virtual total
    r1 = gen_rows(wl, 50_000, 1000, 3000)
    r2 = gen_rows(wl, 50_000_000, 1000, 3000)
    assert np.array_equal(r1[0], r2[0]) and np.array_equal(r1[1], r2[1])


def test_run_digest_is_order_and_content_sensitive():
    from fastapriori_b200 import utils as Utils
    levels = {1: (np.arange(3, dtype=np.int32).reshape(-1, 1), np.array([9, 8, 7], np.int32)),
              2: (np.array([[0, 1], [0, 2]], np.int32), np.array([5, 4], np.int32)), 3: (np.zeros((0, 3), np.int32), np.zeros(0, np.int32))}
    d = Utils.runDigest(levels, ["a", "b", "c"], np.array([1, -1, 0]), (2, 3, 4, 5))
    assert Utils.recommendsText(np.array([1, -1, 0]), ["a", "b", "c"]) == b"b\n0\na\n"
    assert d == Utils.runDigest({k: levels[k] for k in (2, 1, 3)}, ["a", "b", "c"], b"b\n0\na\n", (2, 3, 4, 5))   # empty levels and dict order do not matter
    bumped = {1: levels[1], 2: (levels[2][0], np.array([5, 3], np.int32))}
    assert Utils.runDigest(bumped, ["a", "b", "c"], np.array([1, -1, 0]), (2, 3, 4, 5))["itemsets"] != d["itemsets"]
    assert Utils.runDigest(levels, ["a", "b", "c"], np.array([1, 0, 0]), (2, 3, 4, 5))["recommends"] != d["recommends"]
    assert Utils.runDigest(levels, ["a", "b", "c"], np.array([1, -1, 0]), (2, 3, 4, 6))["all"] != d["all"]


def test_golden_digests_cover_the_baseline_configs():
    import json
    g = json.load(open(os.path.join(ROOT, "tests", "golden", "digests.json")))
    for key in ("c1:100000:10000", "c2:1000000:100000", "c4:4000000:20000", "c5:2000000:40000", "c4:100000000:100000"):
        assert key in g and len(g[key]["all"]) == 32 and g[key]["itemsets_n"] > 0


def test_levels_of_a_closed_engine_do_not_need_the_context():
    from fastapriori_b200.fast_apriori import FreqItemsets

    class Closed:
        _ctx = None

        def wait_copies(self):
            raise AssertionError("must not be called on a closed engine")

    f = FreqItemsets({1: (np.zeros((1, 1), np.int32), np.ones(1, np.int32))}, Closed(), pending=True)
    assert len(f) == 1
