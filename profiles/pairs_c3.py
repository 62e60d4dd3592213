"""K3 at BASELINE config-3 shape: dense 2-itemset stress, N transactions x 16384 items, pair counts by the tcgen05
int8 GEMM.  The CSR is synthesised on the device (Bernoulli rows at ~`density`), attached through the C-ABI, and
only K1/K2/K3 run.  Prints one JSON line with times and int8 op rates.

usage: python profiles/pairs_c3.py [--rows 10000000] [--items 16384] [--density 0.02] [--impl tc|popc] [--check]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch

from fastapriori_b200 import MiningContext


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--items", type=int, default=16384)
    ap.add_argument("--density", type=float, default=0.02)
    ap.add_argument("--impl", default="tc")
    ap.add_argument("--check", action="store_true", help="recount a sample of pairs with torch")
    ap.add_argument("--repeat", type=int, default=2)
    a = ap.parse_args()
    os.environ["FA_PAIR_IMPL"] = a.impl
    sc = MiningContext(device=0)
    eng = sc.engine
    per_row = max(2, int(a.items * a.density))
    g = torch.Generator(device="cuda").manual_seed(3)
    # fixed row length keeps the generator trivial; duplicates inside a row are legal (they collapse in the bitmap)
    ids = torch.randint(0, a.items, (a.rows * per_row,), dtype=torch.int32, device="cuda", generator=g)
    off = torch.arange(0, a.rows + 1, dtype=torch.int64, device="cuda") * per_row
    torch.cuda.synchronize()
    res = {}
    for rep in range(a.repeat):
        eng.attach_transactions(off.data_ptr(), ids.data_ptr(), a.rows, a.items)
        counts = eng.count_items()
        order = np.argsort(-counts.astype(np.int64), kind="stable").astype(np.int32)
        id_to_rank = np.empty(a.items, dtype=np.int32)
        id_to_rank[order] = np.arange(a.items, dtype=np.int32)
        eng.build_bitmaps(id_to_rank, counts[order])
        eng.reset_stats()
        t0 = time.time()
        min_count = int(a.rows * a.density * a.density * 1.05)  # a little above the expected pair support
        n2 = eng.count_pairs(min_count)
        wall = time.time() - t0
        st = eng.stats()
        res = {"rows": a.rows, "items": a.items, "nnz": int(ids.numel()), "impl": a.impl, "frequent_pairs": int(n2),
               "ms_pairs_kernel": st["ms_pairs"], "ms_threshold": st["ms_threshold"], "wall_s_count_pairs": wall,
               "ops_2NF2": st["pair_ops"], "int8_TOPs_nominal": st["pair_ops"] / (st["ms_pairs"] * 1e-3) / 1e12,
               "int8_TOPs_computed_upper_triangle": st["pair_ops"] * (0.5 + 64.0 / a.items) / (st["ms_pairs"] * 1e-3) / 1e12,
               "bitmap_GB": st["bitmap_bytes"] / 1e9}
        res["frac_of_measured_int8_peak_3237"] = res["int8_TOPs_computed_upper_triangle"] / 3237.0
    if a.check:
        sets2, counts2 = eng.get_level(2)
        rng = np.random.default_rng(0)
        rows = torch.arange(a.rows, device="cuda").repeat_interleave(per_row)
        rk = torch.from_numpy(id_to_rank).cuda()[ids.long()]
        for i in rng.choice(len(counts2), size=min(5, len(counts2)), replace=False).tolist():
            x, y = sets2[i].tolist()
            hx = torch.zeros(a.rows, dtype=torch.bool, device="cuda"); hx[rows[rk == x]] = True
            hy = torch.zeros(a.rows, dtype=torch.bool, device="cuda"); hy[rows[rk == y]] = True
            assert int((hx & hy).sum()) == int(counts2[i]), (x, y)
        res["checked_pairs"] = "ok"
    print(json.dumps(res), flush=True)
    sc.close()


if __name__ == "__main__":
    main()
