#!/bin/bash
# Round-2 GPU session O (1 GPU): CTA-pair K3 (k_pair_tc4) as the default above 1024 items; suite; default bench; c2 with p256 forced.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1200 python -m pytest tests -m gpu -q > $O/o_pytest.log 2>&1; echo "rc=$?" >> $O/o_pytest.log; tail -6 $O/o_pytest.log
echo "== bench default"; timeout 900 python bench.py > $O/o_bench.json 2> $O/o_bench.err; echo "rc=$?"; tail -c 600 $O/o_bench.err
echo "== c2 headline with the pair kernel forced"; FA_PAIR_TILE=p256 timeout 300 python bench.py --extras none --no-cpu-baseline --no-digest > $O/o_bench_c2_p256.json 2>/dev/null; echo "rc=$?"
echo "== smoke"; python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
ls -la $O | grep " o_"
