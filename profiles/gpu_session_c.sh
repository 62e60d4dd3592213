#!/bin/bash
# Round-2 GPU session C: full GPU suite, K3 128x128 vs 256x128 tile, default bench line, ncu of the 256x128 kernel.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1200 python -m pytest tests -m gpu -q > $O/c_pytest.log 2>&1; echo "rc=$?" >> $O/c_pytest.log; tail -8 $O/c_pytest.log
echo "== K3 A/B at 1 M x 16384"
FA_PAIR_TILE=128 python profiles/pairs_c3.py --rows 1000000 --repeat 2 > $O/c_pairs_128.json 2>&1; cat $O/c_pairs_128.json | cut -c1-400
FA_PAIR_TILE=256 python profiles/pairs_c3.py --rows 1000000 --repeat 2 --check > $O/c_pairs_256.json 2>&1; cat $O/c_pairs_256.json | cut -c1-400
echo "== bench default"; timeout 900 python bench.py > $O/c_bench.json 2> $O/c_bench.err; echo "rc=$?"; tail -c 600 $O/c_bench.err
FA_PAIR_TILE=128 timeout 900 python bench.py --no-cpu-baseline > $O/c_bench_tile128.json 2> $O/c_bench_tile128.err; echo "rc=$?"
echo "== ncu full, K3 256x128"
timeout 600 ncu --set full --clock-control none -k regex:k_pair_tc2 -c 1 -o $O/c_k3_256 python profiles/pairs_c3.py --rows 1000000 --repeat 1 > $O/c_ncu_k3.log 2>&1; echo "rc=$?"
ls -la $O | grep " c_"
