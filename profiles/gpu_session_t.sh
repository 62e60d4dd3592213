#!/bin/bash
# Round-2 GPU session T (1 GPU): A/B of the level loop, K5 totals read back (FA_K5_WAIT=1, exact sizes) vs left on the device (default), same box.
set -u
O=gpurun_out
mkdir -p $O
for rep in 1 2; do for w in 0 1; do
FA_K5_WAIT=$w timeout 600 python bench.py --extras none --no-cpu-baseline --no-digest --steps 10 > $O/t_bench_wait${w}_$rep.json 2> $O/t_bench.err; echo "rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/t_bench_wait${w}_$rep.json"))
print("wait=$w rep=$rep ms_per_step", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), {k:round(v,2) for k,v in d["device_ms_per_step"].items()}, d["gpu_launches"])
PY
done; done
FA_DEBUG_LEVELS=1 timeout 300 python bench.py --extras none --no-cpu-baseline --no-digest --skip-e2e --steps 1 --warmup 0 2>&1 >/dev/null | grep "^\[level" | head -30 > $O/t_levels.log; cat $O/t_levels.log
