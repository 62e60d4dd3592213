"""Generates the mid-size golden fixtures under tests/golden/quest_*/ with the CPU oracle.

The reference cannot run here (no JVM) and ships no fixtures, so these vectors are produced by oracle/fa_oracle.cpp —
the restatement that is itself pinned by the hand-derived appendix_b vector and by the brute-force checker.  They
freeze its output: the CPU suite checks the oracle against them (regressions in the oracle), the GPU suite checks the
CUDA path against the same bytes without needing the oracle at all.

    python tests/golden/make_quest_golden.py        # rewrites the fixtures (deterministic)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from fastapriori_b200 import quest  # noqa: E402
from oracle import oracle_api  # noqa: E402

CASES = {
    # name: (rows, avg_len, avg_pat, n_pat, n_items, seed, users, min_support)
    "quest_t8": (3000, 8, 4, 25, 60, 101, 120, 0.03),
    "quest_t12_deep": (2500, 12, 6, 15, 40, 202, 150, 0.05),
}


def main():
    for name, (rows, T, I, L, N, seed, users, sup) in CASES.items():
        d = os.path.join(ROOT, "tests", "golden", name)
        os.makedirs(d, exist_ok=True)
        off, items = quest.generate(rows, T, I, L, N, seed)
        uoff, uitems = quest.truncate_rows(*quest.generate(users, T, I, L, N, seed + 1), seed=seed + 2)
        dtxt, utxt = quest.to_text(off, items), quest.to_text(uoff, uitems)
        r = oracle_api.run(dtxt, utxt, sup)
        assert r.status == 0 and not r.level1_ties, "fixtures must not depend on the tie rule"
        open(os.path.join(d, "D.dat"), "wb").write(dtxt)
        open(os.path.join(d, "U.dat"), "wb").write(utxt)
        open(os.path.join(d, "min_support.txt"), "w").write(repr(sup) + "\n")
        open(os.path.join(d, "freqItemset.part-00000"), "wb").write(r.freqitemset_text)
        open(os.path.join(d, "recommends.part-00000"), "wb").write(r.recommends_text)
        counts = sorted(" ".join(r.freq_items[x] for x in sorted(s, reverse=True)) + f"[{c}]" for s, c in r.itemset_dict().items())
        open(os.path.join(d, "counts.txt"), "w").write("\n".join(counts) + "\n")
        ant, cons, conf = r.rules
        with open(os.path.join(d, "rules.txt"), "w") as f:
            for a, c, cf in zip(ant.tolist(), cons.tolist(), conf.tolist()):
                f.write(" ".join(r.freq_items[x] for x in sorted((x for x in a if x >= 0), reverse=True)) + "|" + r.freq_items[c] + "|" + float(cf).hex() + "\n")
        print(name, "rows", rows, "itemsets", len(counts), "max_k", max(k for k, (s, _) in r.levels.items() if len(s)), "rules", len(cons),
              "bytes", sum(os.path.getsize(os.path.join(d, f)) for f in os.listdir(d)))


if __name__ == "__main__":
    main()
