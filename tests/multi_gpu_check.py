"""Run under torchrun with N >= 2 ranks (one per GPU): the row-sharded NCCL path must give, on every rank,
exactly what one GPU gives on the whole dataset — itemsets, supports and recommendations."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from fastapriori_b200 import AssociationRules, FastApriori, MiningContext, TokenizedLines, quest  # noqa: E402


def main():
    import faulthandler
    faulthandler.dump_traceback_later(int(os.environ.get("FA_HANG_DUMP_S", "120")), exit=True)  # a hang must leave a trace, not burn GPU time
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    n_rows, n_items, n_users, sup = 120_000, 300, 4000, 0.01
    off, items = quest.generate(n_rows, 12, 5, 80, n_items, seed=5)
    uoff, uitems = quest.truncate_rows(*quest.generate(n_users, 12, 5, 80, n_items, seed=6), seed=2)
    full = TokenizedLines.from_int_csr(off, items, n_items)
    users = TokenizedLines.from_int_csr(uoff, uitems, n_items)

    results = {}
    for mode in ("native", "callback"):  # the library's own NCCL communicator, then the torch.distributed hook
        sc = MiningContext(device=local, exchange=mode)
        assert sc.world_size == world and sc.exchange == mode
        lo, hi = sc.shard_bounds(n_rows)
        shard = TokenizedLines(full.offsets[lo:hi + 1] - full.offsets[lo], full.ids[full.offsets[lo]:full.offsets[hi]], full.vocab)
        ulo, uhi = sc.shard_bounds(n_users)
        ushard = TokenizedLines(users.offsets[ulo:uhi + 1] - users.offsets[ulo], users.ids[users.offsets[ulo]:users.offsets[uhi]], users.vocab)
        fa = FastApriori(sup)
        fis, i2r, fi = fa.run(sc, shard)
        rec = AssociationRules(fis, fi, i2r).run(sc, ushard, as_arrays=True)
        assert fa.counters["count"] == n_rows
        gathered = [None] * world
        dist.all_gather_object(gathered, rec.tolist())
        results[mode] = (fis.as_dict(), fi, np.array([r for part in gathered for r in part], dtype=np.int32), fis.max_k)
        sc.close()

    single = MiningContext(device=local, distributed=False)
    fa1 = FastApriori(sup)
    fis1, i2r1, fi1 = fa1.run(single, full)
    rec1 = AssociationRules(fis1, fi1, i2r1).run(single, users, as_arrays=True)
    want = fis1.as_dict()
    for mode, (got, fi, rec_all, max_k) in results.items():
        assert fi == fi1, f"{mode}: rank order differs"
        assert got == want, f"{mode}, rank {rank}: sharded itemsets/supports differ from the single-GPU run"
        assert np.array_equal(rec_all, rec1), f"{mode}, rank {rank}: recommendations differ"
        assert max_k >= 4 and len(got) > 1000
    dist.barrier()
    if rank == 0:
        print(f"multi_gpu_check ok: world={world}, {len(want)} itemsets up to k={fis1.max_k}, exchange native+callback, {int((rec1 >= 0).sum())}/{n_users} users matched", flush=True)
    single.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
