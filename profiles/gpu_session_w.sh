#!/bin/bash
# Round-2 GPU session W (1 GPU): K4 register/occupancy variants built with -D flags into scratch_exp/ (16, 20, 24 warps per SM at 2 KB chunks; 24, 32 at 1 KB chunks).
set -u
O=gpurun_out
mkdir -p $O
cp fastapriori_b200/libfastapriori_b200.so /tmp/lib_keep.so
for v in base ws20 ws24 u2ws24 u2ws32 base; do
  cp scratch_exp/lib_$v.so fastapriori_b200/libfastapriori_b200.so
  timeout 300 python bench.py --steps 6 --warmup 3 --extras none --no-cpu-baseline --skip-e2e 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v', round(d['ms_per_step'],2), 'K4', round(d['device_ms_per_step']['ms_count'],2), 'thr', round(d['device_ms_per_step']['ms_threshold'],2), 'golden', d['digest']['matches_golden'])"
done 2>&1 | tee $O/w_sweep.log
cp /tmp/lib_keep.so fastapriori_b200/libfastapriori_b200.so
