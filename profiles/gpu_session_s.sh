#!/bin/bash
# Round-2 GPU session S (1 GPU): level loop with ONE host/device round trip per level (K5 totals stay on the device).
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1500 python -m pytest tests -m gpu -q -x > $O/s_pytest.log 2>&1; echo "rc=$?" >> $O/s_pytest.log; tail -8 $O/s_pytest.log
echo "== bench c2"; timeout 600 python bench.py --extras none --no-cpu-baseline > $O/s_bench_c2.json 2> $O/s_bench.err; echo "rc=$?"; tail -c 300 $O/s_bench.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/s_bench_c2.json"))
print("ms_per_step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], d["device_ms_per_step"], d["digest"]["matches_golden"], d["gpu_launches"])
PY
