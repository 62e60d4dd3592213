#!/bin/bash
# Round-2 GPU session B: full GPU suite (weighted rows, pair-matrix test), K4 tiles (dynamic units) vs groups, f-4 win, ncu of the tile kernel.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1200 python -m pytest tests -m gpu -q > $O/b_pytest.log 2>&1; echo "rc=$?" >> $O/b_pytest.log; tail -8 $O/b_pytest.log
echo "== bench K4 A/B (headline only)"
timeout 300 python bench.py --extras none --no-cpu-baseline > $O/b_bench_tiles.json 2> $O/b_bench_tiles.err; echo "rc=$?"
FA_K4=groups timeout 300 python bench.py --extras none --no-cpu-baseline > $O/b_bench_groups.json 2> $O/b_bench_groups.err; echo "rc=$?"
echo "== f-4: configs[0]-style rows at 10 M, weighted (auto) vs forced off"
timeout 300 python bench.py --workload c1 --rows-per-gpu 10000000 --users-per-gpu 100000 --extras none --no-cpu-baseline > $O/b_c1_10m_weighted.json 2> $O/b_c1_w.err; echo "rc=$?"
FA_WEIGHTED=0 FA_USER_DEDUP=0 timeout 300 python bench.py --workload c1 --rows-per-gpu 10000000 --users-per-gpu 100000 --extras none --no-cpu-baseline > $O/b_c1_10m_plain.json 2> $O/b_c1_p.err; echo "rc=$?"
timeout 300 python bench.py --workload c1 --extras none --no-cpu-baseline > $O/b_c1_weighted.json 2> $O/b_c1_w2.err; echo "rc=$?"
FA_WEIGHTED=0 FA_USER_DEDUP=0 timeout 300 python bench.py --workload c1 --extras none --no-cpu-baseline > $O/b_c1_plain.json 2> $O/b_c1_p2.err; echo "rc=$?"
echo "== ncu full, K4 tiles at level 9"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_count_tiles -s 6 -c 1 -o $O/b_k4_tiles python bench.py --steps 1 --warmup 0 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/b_ncu_full.log 2>&1; echo "rc=$?"
echo "== ncu full, K3 at cfg-3 shape 1 M rows"
timeout 600 ncu --set full --clock-control none -k regex:k_pair_tc -c 1 -o $O/b_k3 python profiles/pairs_c3.py --rows 1000000 --repeat 1 > $O/b_ncu_k3.log 2>&1; echo "rc=$?"
python profiles/pairs_c3.py --rows 1000000 --repeat 2 > $O/b_pairs_c3_1m.json 2>&1
ls -la $O | grep " b_"
