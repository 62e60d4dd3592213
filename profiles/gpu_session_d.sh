#!/bin/bash
# Round-2 GPU session D (1 GPU): K3 256x128 with 4-k-block look-ahead; suite; default bench; launch list + ncu of K3.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1200 python -m pytest tests -m gpu -q > $O/d_pytest.log 2>&1; echo "rc=$?" >> $O/d_pytest.log; tail -6 $O/d_pytest.log
echo "== K3 at 1 M x 16384"
FA_PAIR_TILE=256 python profiles/pairs_c3.py --rows 1000000 --repeat 2 --check > $O/d_pairs_256.json 2>&1; cat $O/d_pairs_256.json | cut -c1-300
echo "== bench default"; timeout 900 python bench.py > $O/d_bench.json 2> $O/d_bench.err; echo "rc=$?"; tail -c 600 $O/d_bench.err
echo "== reference arm"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/d_bench_ref.json 2> $O/d_bench_ref.err; echo "rc=$?"
echo "== ncu launch list (one step incl. warm-up pass)"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $O/d_launches.csv python bench.py --steps 1 --warmup 1 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/d_ncu_list.log 2>&1; echo "rc=$?"
echo "== ncu full, K3 256x128"
timeout 600 ncu --set full --clock-control none -k regex:k_pair_tc2 -c 1 -o $O/d_k3_256 python profiles/pairs_c3.py --rows 1000000 --repeat 1 > $O/d_ncu_k3.log 2>&1; echo "rc=$?"
echo "== ncu full, K4 (group-wise, default) at level 9"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_count_groups2 -s 6 -c 1 -o $O/d_k4_groups python bench.py --steps 1 --warmup 0 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/d_ncu_k4.log 2>&1; echo "rc=$?"
ls -la $O | grep " d_"
