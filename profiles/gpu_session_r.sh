#!/bin/bash
# Round-2 GPU session R (1 GPU): K4 keeps a group's first extension row in registers for the next group (K4_HELD_BIT); A/B with FA_K4_HOLD=0.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest (K4-touching files)"; timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sizes.py tests/test_gpu_weighted.py -m gpu -q > $O/r_pytest.log 2>&1; echo "rc=$?" >> $O/r_pytest.log; tail -4 $O/r_pytest.log
for h in 1 0; do
echo "== bench c2 FA_K4_HOLD=$h"; FA_K4_HOLD=$h timeout 600 python bench.py --extras none --no-cpu-baseline > $O/r_bench_c2_hold$h.json 2> $O/r_bench$h.err; echo "rc=$?"; tail -c 300 $O/r_bench$h.err
python - <<PY
import json
d=json.load(open("gpurun_out/r_bench_c2_hold$h.json"))
print("ms_per_step", d["ms_per_step"], "ms_count", d["device_ms_per_step"]["ms_count"], "frac", d["roofline"]["frac"], "achieved", d["roofline"]["achieved"], d["digest"]["matches_golden"])
PY
done
