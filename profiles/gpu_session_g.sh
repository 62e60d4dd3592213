#!/bin/bash
# Round-2 GPU session G (1 GPU): K2 warp-slice row walk, full-size trim buffers; suite; default bench; early-fetch A/B; host profile.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1200 python -m pytest tests -m gpu -q > $O/g_pytest.log 2>&1; echo "rc=$?" >> $O/g_pytest.log; tail -6 $O/g_pytest.log
echo "== bench default"; timeout 900 python bench.py > $O/g_bench.json 2> $O/g_bench.err; echo "rc=$?"; tail -c 600 $O/g_bench.err
echo "== early fetch A/B (headline only, twice each)"
for i in 1 2; do
  timeout 300 python bench.py --extras none --no-cpu-baseline --no-digest > $O/g_bench_late_$i.json 2>/dev/null
  FA_FETCH_EARLY=1 timeout 300 python bench.py --extras none --no-cpu-baseline --no-digest > $O/g_bench_early_$i.json 2>/dev/null
done
echo "== host profile"; timeout 300 python profiles/prof_host.py > $O/g_prof_host.txt 2>&1; tail -45 $O/g_prof_host.txt
ls -la $O | grep " g_"
