#!/bin/bash
# usage (on the GPU box): build variants with -D flags into scratch_exp/lib_<name>.so, then `bash profiles/variant_sweep.sh <name>...`:
# swaps each library into the box's scratch copy of the package, runs the bench, prints the step and K4 time
for v in "$@"; do
  cp scratch_exp/lib_$v.so fastapriori_b200/libfastapriori_b200.so
  touch fastapriori_b200/libfastapriori_b200.so
  timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --skip-e2e 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v', round(d['ms_per_step'],2), 'count', round(d['device_ms_per_step']['ms_count'],2), 'itemsets', d['config']['itemsets'], 'rules', d['config']['rules'])"
done
