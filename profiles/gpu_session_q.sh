#!/bin/bash
# Round-2 GPU session Q (1 GPU): K4 with one unit per warp, compile-time prefix slot width and byte row addressing; A/B against r2o.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest (K4-touching files)"; timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sizes.py tests/test_gpu_weighted.py -m gpu -q > $O/q_pytest.log 2>&1; echo "rc=$?" >> $O/q_pytest.log; tail -4 $O/q_pytest.log
echo "== bench c2"; timeout 600 python bench.py --extras none --no-cpu-baseline > $O/q_bench_c2.json 2> $O/q_bench.err; echo "rc=$?"; tail -c 300 $O/q_bench.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/q_bench_c2.json"))
print("ms_per_step", d["ms_per_step"], "value", d["value"], "phases", d.get("phases_ms") or d.get("config",{}).get("phases_ms"))
print("roofline", d["roofline"])
PY
