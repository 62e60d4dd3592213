#!/bin/bash
# Round-2 GPU session U (1 GPU): K4 warps per CTA (FA_K4_WARPS = 8, 4, 2, 1), same box.
set -u
O=gpurun_out
mkdir -p $O
for w in 8 4 2 1 8; do
FA_K4_WARPS=$w timeout 600 python bench.py --extras none --no-cpu-baseline --steps 8 > $O/u_bench_w$w.json 2> $O/u_bench.err; echo "rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/u_bench_w$w.json"))
print("warps=$w ms_per_step", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), "K4", round(d["device_ms_per_step"]["ms_count"],2), d["digest"]["matches_golden"])
PY
done
