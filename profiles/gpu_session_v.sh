#!/bin/bash
# Round-2 GPU session V (1 GPU): suite on the 2-warp K4 CTAs; launch list of one mining pass; ncu --set full of K4 at level 9 (source page).
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1500 python -m pytest tests -m gpu -q -x > $O/v_pytest.log 2>&1; echo "rc=$?" >> $O/v_pytest.log; tail -3 $O/v_pytest.log
echo "== bench c2"; timeout 600 python bench.py --extras none --no-cpu-baseline > $O/v_bench_c2.json 2> $O/v_bench.err; echo "rc=$?"; tail -c 300 $O/v_bench.err
echo "== ncu launch list (one step incl. warm-up pass)"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $O/v_launches.csv python bench.py --steps 1 --warmup 1 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/v_ncu_list.log 2>&1; echo "rc=$?"
echo "== ncu full, K4 at level 9"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_count_groups2 -s 6 -c 1 -o $O/v_k4_groups python bench.py --steps 1 --warmup 0 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/v_ncu_k4.log 2>&1; echo "rc=$?"
ls -la $O | grep " v_"
