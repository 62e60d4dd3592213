#!/bin/bash
# Round-2 GPU session K (1 GPU): K9 second pass test; CTA-pair K3 (k_pair_tc3) correctness in its own process, then timing.
set -u
O=gpurun_out
mkdir -p $O
echo "== K9 second pass"; timeout 300 python -m pytest tests/test_gpu_weighted.py -m gpu -q -k "second_pass" > $O/k_pytest_k9.log 2>&1; echo "rc=$?" >> $O/k_pytest_k9.log; tail -4 $O/k_pytest_k9.log
echo "== CTA-pair K3: tc512 against popc / tc / tc256"; FA_TEST_TC512=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "tensor_core" > $O/k_pytest_tc512.log 2>&1; echo "rc=$?" >> $O/k_pytest_tc512.log; tail -12 $O/k_pytest_tc512.log
echo "== CTA-pair K3 timing at 1 M x 16384"
FA_PAIR_TILE=512 timeout 300 python profiles/pairs_c3.py --rows 1000000 --repeat 2 --check > $O/k_pairs_512.json 2>&1; echo "rc=$?"; tail -c 700 $O/k_pairs_512.json
FA_PAIR_TILE=256 timeout 300 python profiles/pairs_c3.py --rows 1000000 --repeat 2 > $O/k_pairs_256.json 2>&1; tail -c 400 $O/k_pairs_256.json
ls -la $O | grep " k_"
