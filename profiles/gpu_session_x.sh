#!/bin/bash
# Round-2 GPU session X (1 GPU): K2 tile size via its shared-memory budget (FA_K2_SMEM_KB = 200, 100, 64, 48): configs[1] and configs[3] at 20 M rows.
set -u
O=gpurun_out
mkdir -p $O
for kb in 200 100 64 48; do
FA_K2_SMEM_KB=$kb timeout 900 python bench.py --extras c4 --extra-scale 0.2 --no-cpu-baseline --skip-e2e --steps 5 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); e=d['extras']['c4']
print('kb=$kb c2 step', round(d['ms_per_step'],2), 'K2', round(d['device_ms_per_step']['ms_bitmap'],3), '| c4x0.2 step', round(e['ms_per_step'],2), 'K2', round(e['device_ms_per_step']['ms_bitmap'],3), 'golden', d['digest']['matches_golden'], e.get('digest',{}).get('matches_golden'))"
done 2>&1 | tee $O/x_k2_sweep.log
