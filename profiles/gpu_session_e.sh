#!/bin/bash
# Round-2 GPU session E (N GPUs, N = first argument, default 2): sharded == single-GPU test, bench line at N with extras, reference arm under torchrun.
set -u
N=${1:-2}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/e${N}_gpus.txt 2>&1; nproc >> $O/e${N}_gpus.txt
echo "== K3 256x128 with the deferred TMEM-store wait: pair tests + timing"
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sizes.py -m gpu -q -k "pair" > $O/e${N}_pytest_pairs.log 2>&1; echo "rc=$?" >> $O/e${N}_pytest_pairs.log; tail -3 $O/e${N}_pytest_pairs.log
FA_PAIR_TILE=256 python profiles/pairs_c3.py --rows 1000000 --repeat 2 --check > $O/e${N}_pairs_256.json 2>&1; cat $O/e${N}_pairs_256.json | cut -c1-300
echo "== sharded == single GPU"; timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q > $O/e${N}_pytest.log 2>&1; echo "rc=$?" >> $O/e${N}_pytest.log; tail -4 $O/e${N}_pytest.log
echo "== bench --gpus $N"
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 > $O/e${N}_bench.json 2> $O/e${N}_bench.err; echo "rc=$?"; tail -c 800 $O/e${N}_bench.err
echo "== reference arm under torchrun"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus $N --steps 1 --warmup 0 > $O/e${N}_bench_ref.json 2> $O/e${N}_bench_ref.err; echo "rc=$?"
ls -la $O | grep " e${N}_"
