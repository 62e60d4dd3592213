"""world_size-2 gloo test of the row-sharded level loop on CPU: each rank mines its
shard, support counts are summed through MiningContext's all-reduce, and every rank must reproduce the
oracle's result on the whole input.  The per-shard counting is done by tests/fake_engine.py (numpy);
the host logic under test — sharding, global count / minCount, rank order, level loop, user sharding and
the gather in saveRecommends — is the product's."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, tmp, d_text, u_text, sup):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from fake_engine import FakeEngine
        from fastapriori_b200 import AssociationRules, FastApriori, MiningContext, TokenizedLines
        from fastapriori_b200 import utils as Utils
        eng = FakeEngine()
        sc = MiningContext(engine=eng)
        eng.allreduce = sc.allreduce_host
        assert (sc.rank, sc.world_size) == (rank, world)
        data = Utils.shard_lines(sc, TokenizedLines.from_text(d_text))
        users = Utils.shard_lines(sc, TokenizedLines.from_text(u_text))
        fa = FastApriori(sup, 0)
        fis, i2r, fi = fa.run(sc, data)
        rec = AssociationRules(fis, fi, i2r).run(sc, users)
        out = os.path.join(tmp, "out_")
        Utils.saveFreqItemset(sc, out, fis, fi)
        Utils.saveRecommends(sc, out, rec)
        np.save(os.path.join(tmp, f"count_{rank}.npy"), np.array([fa.counters["count"], fa.counters["minCount"], len(data)]))
        with open(os.path.join(tmp, f"fis_{rank}.txt"), "wb") as f:
            f.write(Utils.freqItemsetLines(fis, fi, with_count=True))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_row_sharding_matches_oracle(tmp_path):
    sys.path.insert(0, ROOT)
    from oracle import oracle_api
    rng = np.random.default_rng(4)
    rows = [sorted(set(rng.integers(1, 9, size=rng.integers(1, 6)).tolist())) for _ in range(61)]
    users = [sorted(set(rng.integers(1, 10, size=rng.integers(0, 4)).tolist())) for _ in range(13)]
    d = ("\n".join(" ".join(map(str, r)) for r in rows) + "\n").encode()
    u = ("\n".join(" ".join(map(str, r)) for r in users) + "\n").encode()
    sup = 0.1
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path), d, u, sup), nprocs=2, join=True)
    ora = oracle_api.run(d, u, sup)
    c0, c1 = np.load(tmp_path / "count_0.npy"), np.load(tmp_path / "count_1.npy")
    assert c0[0] == c1[0] == ora.n_lines == 61 and c0[1] == c1[1] == ora.min_count
    assert c0[2] + c1[2] == 61 and abs(int(c0[2]) - int(c1[2])) <= 1          # contiguous, balanced shards
    assert open(tmp_path / "fis_0.txt", "rb").read() == open(tmp_path / "fis_1.txt", "rb").read()   # replicated result
    assert open(tmp_path / "out_freqItemset" / "part-00000", "rb").read() == ora.freqitemset_text
    assert open(tmp_path / "out_recommends" / "part-00000", "rb").read() == ora.recommends_text
