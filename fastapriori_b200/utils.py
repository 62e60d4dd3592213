"""Utils — host-side mirror of the reference's text I/O (src/main/scala/com/hazzacheng/AR/Utils.scala).

readAsRDD / saveFreqItemset / saveRecommends keep the reference's names, path-prefix contract
(`path + "D.dat"`, `output + "freqItemset"`) and output layout (a Hadoop-style directory holding
one `part-00000` plus `_SUCCESS`).
"""
from __future__ import annotations

import os

import numpy as np

from .data import TokenizedLines


def readAsRDD(sc, path: str):
    """Utils.scala:19-27.  Every rank reads the files and keeps its contiguous shard of D.dat rows
    (U.dat is sharded the same way); the vocabulary is built from the whole file so dense ids
    agree across ranks."""
    with open(path + "D.dat", "rb") as f:
        data = TokenizedLines.from_text(f.read())
    with open(path + "U.dat", "rb") as f:
        users = TokenizedLines.from_text(f.read())
    return shard_lines(sc, data), shard_lines(sc, users)


def shard_lines(sc, lines: TokenizedLines) -> TokenizedLines:
    if sc.world_size == 1:
        return lines
    lo, hi = sc.shard_bounds(len(lines))
    a, b = lines.offsets[lo], lines.offsets[hi]
    out = TokenizedLines(lines.offsets[lo:hi + 1] - a, lines.ids[a:b], lines.vocab)
    out.index_base = lo
    return out


def _write_part(dirname: str, lines_bytes: bytes):
    # saveAsTextFile semantics: refuse to overwrite (Utils.scala:39,48 -> FileAlreadyExistsException)
    if os.path.exists(dirname):
        raise FileExistsError(f"Output directory {dirname} already exists")
    os.makedirs(dirname)
    with open(os.path.join(dirname, "part-00000"), "wb") as f:
        f.write(lines_bytes)
    open(os.path.join(dirname, "_SUCCESS"), "wb").close()


def freqItemsetLines(freqItemsets, freqItems, with_count: bool = False) -> bytes:
    """Utils.scala:38 (and :60 with counts): items of each set by rank DESCENDING, original id strings joined by a
    space, lines sorted as Java strings (UTF-16 code-unit order).  ASCII item ids (the normal case) are formatted and
    sorted by the native writer (fa_lines_* in libfastapriori_b200.so, multi-threaded); anything else takes the
    Python path below, which orders by UTF-16 code units explicitly."""
    enc = [s.encode("utf-8", "surrogateescape") for s in freqItems]
    ascii_only = all(max(e, default=0) < 0x80 for e in enc)
    if not ascii_only or any(b"\x00" in e or e.endswith(b" ") for e in enc):
        return _freqItemsetLines_slow(freqItemsets, enc, with_count)
    import ctypes
    from . import _native
    lib = _native.load()
    joined = b"".join(e + b"\n" for e in enc)
    h = ctypes.c_void_p()
    rc = lib.fa_lines_new(joined, len(joined), len(enc), ctypes.byref(h))
    if rc != 0:
        raise _native.NativeError(rc, "fa_lines_new failed")
    try:
        for k in sorted(freqItemsets.levels):
            sets, counts = freqItemsets.levels[k]
            if len(counts) == 0:
                continue
            sets = np.ascontiguousarray(sets, dtype=np.int32)
            counts = np.ascontiguousarray(counts, dtype=np.int32)
            rc = lib.fa_lines_add_level(h, sets.ctypes.data, counts.ctypes.data if with_count else None, len(counts), k)
            if rc != 0:
                raise _native.NativeError(rc, "fa_lines_add_level failed")
        total = ctypes.c_int64()
        lib.fa_lines_finish(h, 1, ctypes.byref(total))
        return ctypes.string_at(lib.fa_lines_data(h), total.value)  # one copy, straight out of the formatter's buffer
    finally:
        lib.fa_lines_free(h)


def _freqItemsetLines_slow(freqItemsets, enc, with_count):
    lines = []
    for k in sorted(freqItemsets.levels):
        sets, counts = freqItemsets.levels[k]
        for row, c in zip(sets.tolist(), counts.tolist()):
            s = b" ".join(enc[r] for r in reversed(row))
            if with_count:
                s += b"[" + str(int(c)).encode() + b"]"
            lines.append(s)
    lines.sort(key=lambda b: b.decode("utf-8", "surrogateescape").encode("utf-16-be", "surrogatepass"))
    return b"".join(l + b"\n" for l in lines)


def saveFreqItemset(sc, output: str, freqItemsets, freqItems):
    """Utils.scala:29-41."""
    if sc.rank == 0:
        _write_part(output + "freqItemset", freqItemsetLines(freqItemsets, freqItems))


def saveFreqItemsetWithCount(sc, output: str, freqItemsets, freqItems):
    """Utils.scala:51-63: the `items[count]` dump (unused by the reference's Main; kept as the parity dump)."""
    if sc.rank == 0:
        _write_part(output + "freqItems", freqItemsetLines(freqItemsets, freqItems, with_count=True))


def loadFreqItemsetWithCount(path_or_bytes):
    """Reads an `items[count]` dump (saveFreqItemsetWithCount above, Utils.scala:51-63) back into what
    FastApriori.run returns: (FreqItemsets, itemToRank, freqItems).  The 1-itemset lines give every frequent item and
    its count, hence the rank order (count descending, canonical tie rule); the other lines become rank rows in
    lexicographic order.  AssociationRules accepts the result directly (it installs the levels with fa_put_level), so
    the rule stage can be re-run, or audited, without the transactions."""
    import os
    from .fast_apriori import FreqItemsets, rank_order
    if isinstance(path_or_bytes, (bytes, bytearray)):
        raw = bytes(path_or_bytes)
    else:
        p = path_or_bytes
        if os.path.isdir(p):
            p = os.path.join(p, "part-00000")
        with open(p, "rb") as f:
            raw = f.read()
    by_k = {}
    for ln in raw.split(b"\n"):
        if not ln:
            continue
        body, _, tail = ln.rpartition(b"[")
        if not tail.endswith(b"]") or not body:
            raise ValueError(f"not an items[count] line: {ln[:80]!r}")
        toks = body.split(b" ")
        by_k.setdefault(len(toks), []).append((toks, int(tail[:-1])))
    if 1 not in by_k:
        return FreqItemsets({1: (np.zeros((0, 1), np.int32), np.zeros(0, np.int32))}), {}, []
    vocab = [t[0].decode("utf-8", "surrogateescape") for t, _ in by_k[1]]
    counts1 = np.array([c for _, c in by_k[1]], dtype=np.int32)
    order = rank_order(vocab, counts1, 1)
    freqItems = [vocab[i] for i in order]
    itemToRank = {s: r for r, s in enumerate(freqItems)}
    tok_rank = {s.encode("utf-8", "surrogateescape"): r for s, r in itemToRank.items()}
    levels = {1: (np.arange(len(freqItems), dtype=np.int32).reshape(-1, 1), counts1[order].astype(np.int32))}
    for k in sorted(by_k):
        if k == 1:
            continue
        rows = np.array([[tok_rank[t] for t in toks] for toks, _ in by_k[k]], dtype=np.int32).reshape(-1, k)
        rows.sort(axis=1)
        cnt = np.array([c for _, c in by_k[k]], dtype=np.int32)
        perm = np.lexsort(rows.T[::-1])
        levels[k] = (np.ascontiguousarray(rows[perm]), cnt[perm])
    return FreqItemsets(levels), itemToRank, freqItems


def saveRecommends(sc, output: str, recommends):
    """Utils.scala:43-49: sorted by user index, one item string per line.  With several ranks the
    per-shard results are gathered on rank 0 first."""
    parts = sc.allgather_object(list(recommends))
    if sc.rank == 0:
        allrec = sorted((r for p in parts for r in p), key=lambda x: x[0])
        _write_part(output + "recommends", b"".join(s.encode("utf-8", "surrogateescape") + b"\n" for _, s in allrec))


def rulesLines(sc, rules) -> bytes:
    """Parity dump of the kept rules in the reference's sorted order (the reference never writes its
    rules): `antecedent items (rank descending, like saveFreqItemset)|consequent|confidence as a C99 hex double`.
    `rules` is the AssociationRules object whose run() produced them."""
    n = rules.counters.get("rules", 0)
    ant, cons, conf = sc.engine.get_rules(1, n)
    items = rules.freqItems
    out = []
    for a, c, f in zip(ant.tolist(), cons.tolist(), conf.tolist()):
        out.append(" ".join(items[x] for x in sorted((x for x in a if x >= 0), reverse=True)) + "|" + items[c] + "|" + float(f).hex())
    return ("\n".join(out) + "\n").encode("utf-8", "surrogateescape") if out else b""


def saveRules(sc, output: str, rules):
    if sc.rank == 0:
        _write_part(output + "rules", rulesLines(sc, rules))


# ---- result digests (audit / parity at sizes where comparing the outputs themselves is impractical) -----------------
def levelsDigest(levels) -> str:
    """SHA-256 over every non-empty level in ascending k: (k, n) then the rank rows (int32, lexicographic order) and the
    supports (int32).  Two runs agree on it iff they found the same frequent itemsets with the same supports (and the
    same rank order, which the canonical tie rule fixes)."""
    import hashlib
    import struct
    h = hashlib.sha256()
    for k in sorted(levels):
        sets, counts = levels[k]
        if len(counts) == 0:
            continue
        h.update(struct.pack("<iq", int(k), len(counts)))
        h.update(np.ascontiguousarray(sets, dtype="<i4").tobytes())
        h.update(np.ascontiguousarray(counts, dtype="<i4").tobytes())
    return h.hexdigest()


def recommendsText(rec_ranks, freqItems) -> bytes:
    """The bytes of recommends/part-00000 (Utils.scala:48) for consequent ranks in user order (-1 -> "0")."""
    table = np.array(list(freqItems) + ["0"], dtype=object)
    rec = np.asarray(rec_ranks, dtype=np.int64)
    if len(rec) == 0:
        return b""
    return ("\n".join(table[rec].tolist()) + "\n").encode("utf-8", "surrogateescape")


def runDigest(levels, freqItems, rec_ranks_or_text, rules_digest) -> dict:
    """Digest of one mining run: itemsets + supports, rules (the four words of fa_rules_digest) and recommendations."""
    import hashlib
    rec_text = rec_ranks_or_text if isinstance(rec_ranks_or_text, (bytes, bytearray)) else recommendsText(rec_ranks_or_text, freqItems)
    d = {"itemsets": levelsDigest(levels),
         "rules": ":".join(f"{int(x):x}" for x in rules_digest),
         "recommends": hashlib.sha256(rec_text).hexdigest()}
    d["all"] = hashlib.sha256("|".join(d[k] for k in ("itemsets", "rules", "recommends")).encode()).hexdigest()[:32]
    return d
