import sys, os
sys.path.insert(0, os.getcwd())
os.environ["FA_DEBUG_LEVELS"] = "1"; os.environ["FA_DEBUG_TRIM"] = "1"
from fastapriori_b200 import FastApriori, MiningContext, TokenizedLines, quest
off, items = quest.generate(1000000, 40, 10, 2000, 1000, seed=2)
d = TokenizedLines.from_int_csr(off, items, 1000)
sc = MiningContext(device=0)
for i in range(2):
    if i == 1: print("==== second run", file=sys.stderr, flush=True)
    fa = FastApriori(0.005); fis, _, fi = fa.run(sc, d); print(len(fis), fa.counters)
