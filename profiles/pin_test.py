import sys, time, os
sys.path.insert(0, os.getcwd())
import torch, numpy as np
from fastapriori_b200 import FastApriori, MiningContext, TokenizedLines, quest
torch.cuda.init()
t=time.time(); x=torch.empty(32*1024*1024, dtype=torch.int32, pin_memory=True); print('pin 128MB', round(time.time()-t,3)); 
t=time.time(); y=torch.empty(32*1024*1024, dtype=torch.int32, pin_memory=True); print('pin 128MB again', round(time.time()-t,3))
del x, y
t=time.time(); y=torch.empty(32*1024*1024, dtype=torch.int32, pin_memory=True); print('pin 128MB cached', round(time.time()-t,3))
off, items = quest.generate(1000000, 40, 10, 2000, 1000, seed=2)
data = TokenizedLines.from_int_csr(off, items, 1000)
sc = MiningContext(device=0)
keep=[]
for rep in range(4):
    t=time.time(); fis,_,_ = FastApriori(0.005).run(sc, data); t1=time.time()-t; _=fis.levels; t2=time.time()-t
    print('rep',rep,'run',round(t1,3),'with levels',round(t2,3), 'keeping previous result' if rep<2 else 'dropping previous')
    if rep<2: keep.append(fis)
    else: keep.clear(); del fis
for rep in range(2):
    t=time.time(); fis,_,_ = FastApriori(0.005).run(sc, data, fetch=False); t1=time.time()-t
    lv={k: sc.engine.get_level(k) for k in range(2, 22)}; t2=time.time()-t
    print('pageable rep',rep,'run',round(t1,3),'with levels',round(t2,3))
