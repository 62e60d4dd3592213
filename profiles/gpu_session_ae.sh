#!/bin/bash
# Round-2 GPU session AE (1 GPU): DRAM / L2 traffic of every K4 launch of one pass with the final kernel (configs[1]; configs[3] shape at 20 M rows).
set -u
O=gpurun_out
mkdir -p $O
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum
timeout 600 ncu --metrics $M --clock-control none -k regex:k_count_groups2 -c 19 --csv --log-file $O/ae_k4_traffic_c2.csv python bench.py --steps 1 --warmup 0 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/ae_ncu_traffic.log 2>&1; echo "rc=$?"
timeout 600 ncu --metrics $M --clock-control none -k regex:k_count_groups2 -c 12 --csv --log-file $O/ae_k4_traffic_c4.csv python bench.py --workload c4 --rows-per-gpu 20000000 --users-per-gpu 20000 --steps 1 --warmup 0 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/ae_ncu_traffic_c4.log 2>&1; echo "rc=$?"
python profiles/k4_traffic.py $O/ae_k4_traffic_c2.csv $O/ae_k4_traffic_c4.csv; cp profiles/k4_traffic.json $O/ae_k4_traffic.json
