#!/bin/bash
# Round-2 GPU session I (1 GPU): trim kernel over item slices, K3 back to one B group; suite; default bench; launch list.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1200 python -m pytest tests -m gpu -q > $O/i_pytest.log 2>&1; echo "rc=$?" >> $O/i_pytest.log; tail -6 $O/i_pytest.log
echo "== bench default"; timeout 900 python bench.py > $O/i_bench.json 2> $O/i_bench.err; echo "rc=$?"; tail -c 600 $O/i_bench.err
echo "== ncu launch list"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $O/i_launches.csv python bench.py --steps 1 --warmup 1 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/i_ncu_list.log 2>&1; echo "rc=$?"
ls -la $O | grep " i_"
