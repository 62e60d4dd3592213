"""The five BASELINE.json configurations as parameters of the seeded Quest generator (quest.py): one table shared
by bench.py, the parity tests and the script that freezes the golden digests (tests/golden/make_digests.py).
Workload tooling, not part of the mining path."""
from __future__ import annotations

import numpy as np

WORKLOADS = {
    # configs[0]: the reference's default min-support on T10I4D100K-style rows; few patterns so that pairs clear 9.2 %
    "c1": dict(name="T10I4D100K", cfg="configs[0]", avg_len=10.0, avg_pat=4.0, n_pat=30, n_items=1000, min_support=0.092, seed=1,
               rows=100_000, users=10_000),
    # configs[1]: the headline (bitmap AND + popcount path, deep lattice)
    "c2": dict(name="T40I10D1M", cfg="configs[1]", avg_len=40.0, avg_pat=10.0, n_pat=2000, n_items=1000, min_support=0.005, seed=2,
               rows=1_000_000, users=100_000),
    # configs[3]: the scaling workload, row-sharded over the GPUs of a node
    "c4": dict(name="T20I6D100M", cfg="configs[3]", avg_len=20.0, avg_pat=6.0, n_pat=2000, n_items=10000, min_support=0.005, seed=4,
               rows=100_000_000, users=100_000),
    # configs[4]: the full rule + recommendation pipeline at low support
    "c5": dict(name="T20I6D50M+1M users", cfg="configs[4]", avg_len=20.0, avg_pat=6.0, n_pat=2000, n_items=10000, min_support=0.0005,
               seed=5, rows=50_000_000, users=1_000_000),
}


def gen_rows(wl: dict, total_rows: int, lo: int, hi: int):
    """Rows [lo, hi) of the workload's virtual dataset of `total_rows` rows: (offsets int64, items int32 in 1..n_items)."""
    from . import quest
    return quest.generate(total_rows, wl["avg_len"], wl["avg_pat"], wl["n_pat"], wl["n_items"], wl["seed"], lo, hi)


def gen_users(wl: dict, total_users: int, lo: int, hi: int):
    """User baskets [lo, hi): rows of the same generator under another seed, each cut to a random non-empty prefix.
    The cut of user u depends on u alone (not on the shard), so any sharding yields the same baskets."""
    from . import quest
    total_users = max(total_users, 1)
    off, items = quest.generate(total_users, wl["avg_len"], wl["avg_pat"], wl["n_pat"], wl["n_items"], wl["seed"] + 1000, lo, hi)
    lens = np.diff(off)
    u = np.arange(lo, hi, dtype=np.uint64)
    # splitmix64 of (user index, seed) -> uniform in [0, 1)
    z = (u + np.uint64(wl["seed"])) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(0x1234567)
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    z = z ^ (z >> np.uint64(31))
    frac = (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    keep = np.minimum(lens, 1 + (frac * lens).astype(np.int64))
    new_off = np.zeros(len(off), dtype=np.int64)
    np.cumsum(keep, out=new_off[1:])
    idx = np.repeat(off[:-1] - new_off[:-1], keep) + np.arange(int(new_off[-1]), dtype=np.int64)
    return new_off, items[idx]
