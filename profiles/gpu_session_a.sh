#!/bin/bash
# Round-2 GPU session A: full GPU test suite, default bench line (with extras), reference arm, K4 A/B, ncu evidence.
# usage (from the repo root, under gpurun): bash profiles/gpu_session_a.sh
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > $O/a_gpu.txt 2>&1
nproc > $O/a_nproc.txt; free -g >> $O/a_nproc.txt
echo "== pytest (tiles)"; timeout 900 python -m pytest tests -m gpu -q -x > $O/a_pytest.log 2>&1; echo "rc=$?" >> $O/a_pytest.log; tail -5 $O/a_pytest.log
if ! grep -q "rc=0" $O/a_pytest.log; then
  echo "== pytest (group-wise K4)"; FA_K4=groups timeout 900 python -m pytest tests -m gpu -q > $O/a_pytest_groups.log 2>&1; echo "rc=$?" >> $O/a_pytest_groups.log; tail -5 $O/a_pytest_groups.log
fi
echo "== bench K4 A/B (headline only)"
timeout 300 python bench.py --extras none --no-cpu-baseline > $O/a_bench_tiles.json 2> $O/a_bench_tiles.err; echo "rc=$?"
FA_K4=groups timeout 300 python bench.py --extras none --no-cpu-baseline > $O/a_bench_groups.json 2> $O/a_bench_groups.err; echo "rc=$?"
echo "== bench default"; timeout 900 python bench.py > $O/a_bench.json 2> $O/a_bench.err; echo "rc=$?"; tail -c 600 $O/a_bench.err
echo "== reference arm"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/a_bench_ref.json 2> $O/a_bench_ref.err; echo "rc=$?"
echo "== ncu launch list"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $O/a_launches.csv python bench.py --steps 1 --warmup 1 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/a_ncu_list.log 2>&1; echo "rc=$?"
echo "== ncu K4 dram traffic, all launches of one step (c2)"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum --clock-control none -k regex:k_count_tiles -c 19 --csv --log-file $O/a_k4_traffic_c2.csv python bench.py --steps 1 --warmup 0 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/a_ncu_traffic.log 2>&1; echo "rc=$?"
echo "== ncu K4 dram traffic, cfg-4 shape 20 M rows (database >> L2)"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum --clock-control none -k regex:k_count_tiles -c 12 --csv --log-file $O/a_k4_traffic_c4.csv python bench.py --workload c4 --rows-per-gpu 20000000 --users-per-gpu 20000 --steps 1 --warmup 0 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/a_ncu_traffic_c4.log 2>&1; echo "rc=$?"
echo "== ncu full, K4 at a heavy level"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_count_tiles -s 6 -c 1 -o $O/a_k4_tiles python bench.py --steps 1 --warmup 0 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/a_ncu_full.log 2>&1; echo "rc=$?"
ls -la $O | tail -30
