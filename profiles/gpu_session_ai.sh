#!/bin/bash
# Round-2 GPU session AI (1 GPU): 16-bit ids on the host-buffer path (fa_load_transactions_u16): test + headline e2e.
set -u
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "compact or appendix or waited or random_small" > $O/ai_pytest.log 2>&1; echo "rc=$?" >> $O/ai_pytest.log; tail -3 $O/ai_pytest.log
timeout 600 python bench.py --extras none --no-cpu-baseline > $O/ai_bench.json 2> $O/ai_bench.err; echo "rc=$?"; tail -c 300 $O/ai_bench.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/ai_bench.json") if l.startswith("{")][-1])
print("ms_per_step", round(d["ms_per_step"],2), "value", round(d["value"]/1e6,2), "e2e", d["e2e"], d["digest"]["matches_golden"])
PY
