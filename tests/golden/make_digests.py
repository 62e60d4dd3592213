"""Freezes the digests the CPU oracle produces for the BASELINE.json workloads (same seeds as bench.py):
tests/golden/digests.json, keyed "<workload>:<rows_total>:<users_total>".

Run where the oracle has the time and the memory (the CPU container: `python tests/golden/make_digests.py c1 c2 ...`);
the GPU tests (tests/test_gpu_sizes.py) and bench.py only read the JSON.  The oracle runs in its packed, no-dedup mode
with hashed rule lookups — the modes tests/test_oracle.py holds to the naive restatement, output for output.
The digest itself is fastapriori_b200.utils.runDigest: SHA-256 over every level's rank rows + supports, the four words
of the rule digest (kept rules as a multiset, and their sorted (confidence, item id) sequence), SHA-256 of the
recommends text.

usage: make_digests.py <case> [<case> ...]   cases: c1 c2 c2x2 c2x4 c2x8 c4s c5s c4 c5 (see CASES)
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from fastapriori_b200 import utils as Utils  # noqa: E402
from fastapriori_b200.workloads import WORKLOADS, gen_rows, gen_users  # noqa: E402
from oracle import oracle_api  # noqa: E402

# name -> (workload, rows, users)
CASES = {
    "c1": ("c1", 100_000, 10_000),
    "c2": ("c2", 1_000_000, 100_000),
    "c2x2": ("c2", 2_000_000, 200_000),   # bench.py --gpus 2/4/8: weak scaling of the headline
    "c2x4": ("c2", 4_000_000, 400_000),
    "c2x8": ("c2", 8_000_000, 800_000),
    "c4s": ("c4", 4_000_000, 20_000),     # cfg-4 / cfg-5 shape at a size the GPU test suite runs in seconds
    "c5s": ("c5", 2_000_000, 40_000),
    "c4": ("c4", 100_000_000, 100_000),   # the full sizes bench.py reports under "extras"
    "c5": ("c5", 50_000_000, 1_000_000),
}
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "digests.json")


def main():
    names = sys.argv[1:] or ["c1"]
    for name in names:
        key, rows, users = CASES[name]
        wl = WORKLOADS[key]
        t0 = time.time()
        off, items = gen_rows(wl, rows, 0, rows)
        uoff, uitems = gen_users(wl, users, 0, users)
        r = oracle_api.run_csr(off, items, uoff, uitems, wl["min_support"], fast_rules=True, fast_cands=True, packed=True,
                               no_dedup=True, light=True)
        dig = Utils.runDigest(r.levels, r.freq_items, r.recommends_text, r.rules_digest)
        dig.update(itemsets_n=sum(len(c) for _, c in r.levels.values()), max_k=max((k for k, (_, c) in r.levels.items() if len(c)), default=0),
                   frequent_items=len(r.freq_items), raw_rules=r.n_raw, rules_n=r.n_kept, min_count=r.min_count,
                   oracle_seconds=round(time.time() - t0, 1), oracle_times={k: round(v, 2) for k, v in r.times.items()})
        try:
            allv = json.load(open(OUT))
        except Exception:
            allv = {}
        allv[f"{key}:{rows}:{users}"] = dig
        json.dump(allv, open(OUT, "w"), indent=1, sort_keys=True)
        print(name, json.dumps(dig), flush=True)


if __name__ == "__main__":
    main()
