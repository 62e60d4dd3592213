#!/bin/bash
# Round-2 GPU session AF (1 GPU): host-side profile of a configs[3] step at the per-rank size of an 8-GPU run (12.5 M rows, 12.5 k users).
set -u
O=gpurun_out
mkdir -p $O
timeout 600 python profiles/prof_host_c4.py > $O/af_prof_host_c4.txt 2>&1; echo "rc=$?"; head -70 $O/af_prof_host_c4.txt
