"""Independent brute-force checker for tiny inputs (test infrastructure, see fa_oracle.cpp).

It does not follow the reference's algorithm: supports come from enumerating every subset
of every transaction, rules and the prune come straight from their definitions
(SURVEY.md Appendix A).  It exists to pin oracle/fa_oracle.cpp on random small inputs,
because the reference itself offers no golden vectors.  Pure Python; only for rows of
<= ~12 distinct items.
"""
from __future__ import annotations

import math
from itertools import combinations


def _is_int(s: str) -> bool:
    t = s[1:] if s[:1] in "+-" else s
    return t.isdigit() and t.isascii() and -2**31 <= int(s) <= 2**31 - 1


def tokenize(text: str):
    """Utils.scala:21 — records split on \\n / \\r\\n / \\r, Java trim, split on \\s+."""
    if not text:
        return []
    lines = text.replace("\r\n", "\n").replace("\r", "\n").split("\n")
    if lines and lines[-1] == "":
        lines.pop()
    out = []
    for ln in lines:
        ln = ln.strip("".join(chr(c) for c in range(0x21)))
        toks = [t for t in __import__("re").split(r"[ \t\n\x0b\f\r]+", ln)] if ln else [""]
        out.append(toks)
    return out


def mine(d_text: str, u_text: str, min_support: float):
    data = tokenize(d_text)
    users = tokenize(u_text)
    n = len(data)
    min_count = int(math.ceil(min_support * n))
    occ = {}
    for line in data:
        for t in line:
            occ[t] = occ.get(t, 0) + 1
    freq = [t for t, c in occ.items() if c >= min_count]
    # canonical tie rule shared with the oracle: count desc, ints by value, then byte order
    freq.sort(key=lambda t: (-occ[t], 0 if _is_int(t) else 1, int(t) if _is_int(t) else 0, t.encode()))
    rank = {t: i for i, t in enumerate(freq)}
    support = {}
    for line in data:
        s = sorted({rank[t] for t in line if t in rank})
        for k in range(2, len(s) + 1):
            for sub in combinations(s, k):
                support[sub] = support.get(sub, 0) + 1
    itemsets = {(rank[t],): occ[t] for t in freq}
    itemsets.update({s: c for s, c in support.items() if c >= min_count})
    # rules by definition (A-10), prune by definition (A-11)
    conf = {}
    for s, c in itemsets.items():
        if len(s) >= 2:
            for i in s:
                ant = tuple(x for x in s if x != i)
                conf[(ant, i)] = c / itemsets[ant]
    kept = {}
    for (ant, cons) in sorted(conf, key=lambda r: len(r[0])):
        cf = conf[(ant, cons)]
        if len(ant) == 1:
            kept[(ant, cons)] = cf
            continue
        ok = True
        for a in ant:
            sub = tuple(x for x in ant if x != a)
            if (sub, cons) not in kept or not kept[(sub, cons)] < cf:
                ok = False
                break
        if ok:
            kept[(ant, cons)] = cf
    has_rules = any(len(s) >= 2 for s in itemsets)
    order = sorted(kept, key=lambda r: (-kept[r], int(freq[r[1]]) if _is_int(freq[r[1]]) else 0))
    recs = []
    for line in users:
        u = {rank[t] for t in line if t in rank}
        rec = "0"
        if u:
            for ant, cons in order:
                if len(ant) <= len(u) and cons not in u and set(ant) <= u:
                    rec = freq[cons]
                    break
        recs.append(rec)
    lines = sorted(" ".join(freq[r] for r in sorted(s, reverse=True)).encode() for s in itemsets)
    return {
        "n": n, "min_count": min_count, "freq_items": freq, "itemsets": itemsets, "conf": conf,
        "kept": kept, "order": order, "recommends": recs if has_rules else None,
        "freqitemset_text": b"".join(l + b"\n" for l in lines),
    }
