#!/bin/bash
# Round-2 GPU session H (1 GPU): K3 with two B producer groups, 8-wave splits; c3 recount; suite; default bench; ncu of K3.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1200 python -m pytest tests -m gpu -q > $O/h_pytest.log 2>&1; echo "rc=$?" >> $O/h_pytest.log; tail -6 $O/h_pytest.log
echo "== K3 at 1 M x 16384"
python profiles/pairs_c3.py --rows 1000000 --repeat 2 --check > $O/h_pairs_256.json 2>&1; cat $O/h_pairs_256.json | cut -c1-300
echo "== bench default"; timeout 900 python bench.py > $O/h_bench.json 2> $O/h_bench.err; echo "rc=$?"; tail -c 600 $O/h_bench.err
echo "== ncu full, K3 256x128"
timeout 600 ncu --set full --clock-control none -k regex:k_pair_tc2 -c 1 -o $O/h_k3_256 python profiles/pairs_c3.py --rows 1000000 --repeat 1 > $O/h_ncu_k3.log 2>&1; echo "rc=$?"
ls -la $O | grep " h_"
