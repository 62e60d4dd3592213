#!/bin/bash
# Round-2 GPU session AG (1 GPU): where fa_build_bitmaps spends host time at the configs[3] per-rank shape.
set -u
O=gpurun_out
mkdir -p $O
FA_DEBUG_HOST=1 timeout 600 python bench.py --workload c4 --rows-per-gpu 12500000 --users-per-gpu 12500 --steps 3 --warmup 2 --no-cpu-baseline --skip-e2e --extras none --no-digest 2> $O/ag_host.log > /dev/null; echo "rc=$?"; grep build_bitmaps $O/ag_host.log | tail -12
