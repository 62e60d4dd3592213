"""com.hazzacheng.AR.Main, B200 edition: `python -m fastapriori_b200.main <input> <output> <temp> [minSupport]`.

Same positional contract as the reference (README.md:19-41, Main.scala:24-25): <input> and
<output> are string PREFIXES (`<input>D.dat`, `<input>U.dat`, `<output>freqItemset`,
`<output>recommends`), <temp> is accepted and ignored exactly as the reference ignores it.
minSupport defaults to the reference's hard-coded 0.092 (Main.scala:23); an optional fourth
argument or FA_MIN_SUPPORT overrides it.  Under torchrun each rank takes a row shard.
"""
from __future__ import annotations

import os
import sys
import time

from . import utils as Utils
from .association_rules import AssociationRules
from .context import MiningContext
from .fast_apriori import FastApriori


def main(argv=None):
    args = list(sys.argv[1:] if argv is None else argv)
    input_, output = args[0], args[1]  # IndexError with < 2 args, like the reference's ArrayIndexOutOfBounds
    minSupport = float(args[3]) if len(args) > 3 else float(os.environ.get("FA_MIN_SUPPORT", "0.092"))
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group("nccl")
    sc = MiningContext(verbose=True)
    dataRDD, userRDD = Utils.readAsRDD(sc, input_)

    time1 = time.time()
    freqItemsets, itemToRank, freqItems = FastApriori(minSupport, 500 if os.environ.get("FA_TIE_ORDER") == "spark" else 0,
                                                       tie_order=os.environ.get("FA_TIE_ORDER", "canonical")).run(sc, dataRDD)
    Utils.saveFreqItemset(sc, output, freqItemsets, freqItems)
    sc.log(f"==== Total time for get freqItemsets {int((time.time() - time1) * 1000)}")

    time2 = time.time()
    recommends = AssociationRules(freqItemsets, freqItems, itemToRank, strict=True).run(sc, userRDD)
    Utils.saveRecommends(sc, output, recommends)
    sc.log(f"==== Total time for get recommends {int((time.time() - time2) * 1000)}")
    sc.close()


if __name__ == "__main__":
    main()
