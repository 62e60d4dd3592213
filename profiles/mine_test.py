import sys, time, os
sys.path.insert(0, os.getcwd())
import torch, numpy as np
from fastapriori_b200 import FastApriori, MiningContext, TokenizedLines, quest
off, items = quest.generate(1000000, 40, 10, 2000, 1000, seed=2)
txt = quest.to_text(off[:200001], items[:off[200000]])
d_int = TokenizedLines.from_int_csr(off, items, 1000)
quest.write_text('/tmp/D.dat', off, items)
t=time.time(); d_txt = TokenizedLines.from_text(open('/tmp/D.dat','rb').read()); print('parse', round(time.time()-t,3), len(d_txt), len(d_txt.vocab), d_txt.ids.dtype, d_txt.offsets.dtype, d_txt.ids.flags['C_CONTIGUOUS'])
sc = MiningContext(device=0)
for name, d in (('int', d_int), ('txt', d_txt), ('int', d_int), ('txt', d_txt)):
    fa = FastApriori(0.005)
    t=time.time(); fis,_,fi = fa.run(sc, d); _=fis.levels; dt=time.time()-t
    st = sc.engine.stats(); sc.engine.reset_stats()
    print(name, round(dt,3), 'itemsets', len(fis), 'F', len(fi), {k: round(v,1) for k,v in st.items() if k.startswith('ms_') and v>0.5}, fi[:5])
    del fis
