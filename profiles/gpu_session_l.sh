#!/bin/bash
# Round-2 GPU session L (1 GPU): one mbarrier arrival per producer warp in k_pair_tc2 / k_pair_tc3; correctness, timing, ncu.
set -u
O=gpurun_out
mkdir -p $O
echo "== pair tests (tc512 included)"; FA_TEST_TC512=1 timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sizes.py -m gpu -q -x -k "pair or tensor_core" > $O/l_pytest_pairs.log 2>&1; echo "rc=$?" >> $O/l_pytest_pairs.log; tail -6 $O/l_pytest_pairs.log
echo "== timing at 1 M x 16384"
for t in 256 512; do
  FA_PAIR_TILE=$t timeout 300 python profiles/pairs_c3.py --rows 1000000 --repeat 2 --check > $O/l_pairs_$t.json 2>&1; echo "tile $t rc=$?"; tail -c 420 $O/l_pairs_$t.json; echo
done
echo "== ncu full, CTA-pair kernel"
FA_PAIR_TILE=512 timeout 600 ncu --set full --clock-control none -k regex:k_pair_tc3 -c 1 -o $O/l_k3_512 python profiles/pairs_c3.py --rows 1000000 --repeat 1 > $O/l_ncu_k3.log 2>&1; echo "rc=$?"
ls -la $O | grep " l_"
