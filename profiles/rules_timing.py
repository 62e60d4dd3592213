import sys, time, os
sys.path.insert(0, os.getcwd())
import torch, numpy as np
from fastapriori_b200 import FastApriori, AssociationRules, MiningContext, TokenizedLines, quest
off, items = quest.generate(1000000, 40, 10, 2000, 1000, seed=2)
data = TokenizedLines.from_int_csr(off, items, 1000)
uoff, uitems = quest.truncate_rows(*quest.generate(100000, 40, 10, 2000, 1000, seed=1002), seed=7)
users = TokenizedLines.from_int_csr(uoff, uitems, 1000)
sc = MiningContext(device=0); eng = sc.engine
for rep in range(3):
    fis, i2r, fi = FastApriori(0.005).run(sc, data)
    sync = rep == 2
    if sync: torch.cuda.synchronize(); eng.wait_copies()
    t0=time.perf_counter(); n_raw = eng.gen_rules(); t1=time.perf_counter()
    if sync: torch.cuda.synchronize()
    t1b=time.perf_counter(); n_kept = eng.prune_rules(); t2=time.perf_counter()
    ar = AssociationRules(fis, fi, i2r)
    o, r = ar.removeRedundancy(sc, users); t3=time.perf_counter()
    eng.sort_rules(np.array([int(s) for s in fi], dtype=np.int32)); t4=time.perf_counter()
    rec = eng.recommend(o, r); t5=time.perf_counter()
    _ = fis.levels; t6=time.perf_counter()
    print('rep',rep,'synced' if sync else '', {k: round(v*1e3,2) for k,v in dict(gen_rules=t1-t0, prune=t2-t1b, redundancy=t3-t2, sort=t4-t3, recommend=t5-t4, wait_levels=t6-t5).items()}, n_raw, n_kept)
