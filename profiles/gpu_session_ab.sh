#!/bin/bash
# Round-2 GPU session AB (1 GPU): k_mark_active through a shared-memory bitmap, coalesced k_compact_level; lanes per x in the K5 join (FA_JOIN_LPX).
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1500 python -m pytest tests -m gpu -q -x > $O/ab_pytest.log 2>&1; echo "rc=$?" >> $O/ab_pytest.log; tail -4 $O/ab_pytest.log
for lpx in 0 4 2 0; do
FA_JOIN_LPX=$lpx timeout 600 python bench.py --extras none --no-cpu-baseline --skip-e2e --steps 8 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('lpx=$lpx step', round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['device_ms_per_step'].items() if k in ('ms_candgen','ms_threshold','ms_count')}, d['digest']['matches_golden'])"
done 2>&1 | tee $O/ab_sweep.log
