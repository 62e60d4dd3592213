#!/bin/bash
# Round-2 GPU session Y (N GPUs, default 2): the one-round-trip level loop under NCCL (sharded == single GPU test), bench line at N with extras.
set -u
N=${1:-2}
O=gpurun_out
mkdir -p $O
echo "== sharded == single GPU"; timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q > $O/y${N}_pytest.log 2>&1; echo "rc=$?" >> $O/y${N}_pytest.log; tail -4 $O/y${N}_pytest.log
echo "== bench --gpus $N"
date +%s > $O/y${N}_t0
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 > $O/y${N}_bench.json 2> $O/y${N}_bench.err; echo "rc=$?"; tail -c 600 $O/y${N}_bench.err
echo "bench seconds: $(( $(date +%s) - $(cat $O/y${N}_t0) ))"
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/y${N}_bench.json") if l.startswith("{")][-1])
print("N=$N ms_per_step", round(d["ms_per_step"],2), "value", round(d["value"]/1e6,2), "M tx/s e2e", round(d["e2e"]["value"]/1e6,2), {k:round(v,2) for k,v in d["device_ms_per_step"].items()}, d["digest"])
for k,v in d.get("extras",{}).items(): print(" ", k, round(v.get("ms_per_step",0),2), v.get("digest",{}).get("matches_golden"), v.get("digest",{}).get("ranks_equal"))
PY
