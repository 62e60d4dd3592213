#!/bin/bash
# Round-2 GPU session F (1 GPU): K2 with four load chains, cost-aware trimming; suite, default bench, trim A/B.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1200 python -m pytest tests -m gpu -q > $O/f_pytest.log 2>&1; echo "rc=$?" >> $O/f_pytest.log; tail -6 $O/f_pytest.log
echo "== bench default"; timeout 900 python bench.py > $O/f_bench.json 2> $O/f_bench.err; echo "rc=$?"; tail -c 600 $O/f_bench.err
echo "== bench with unconditional trimming (round-1 rule)"; FA_TRIM_ALWAYS=1 timeout 900 python bench.py --no-cpu-baseline > $O/f_bench_trim_always.json 2> $O/f_bench_trim_always.err; echo "rc=$?"
echo "== smoke"; python -c "import __graft_entry__ as g; g.smoke()" > $O/f_smoke.log 2>&1; echo "rc=$?"; tail -2 $O/f_smoke.log
ls -la $O | grep " f_"
