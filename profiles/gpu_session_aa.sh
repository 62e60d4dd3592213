#!/bin/bash
# Round-2 GPU session AA (1 GPU): K8 confidence sort on the bit range that can differ; suite + headline.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1500 python -m pytest tests -m gpu -q > $O/aa_pytest.log 2>&1; echo "rc=$?" >> $O/aa_pytest.log; tail -4 $O/aa_pytest.log
echo "== bench c2 + c5 (rule-heavy)"; timeout 900 python bench.py --extras c5 --no-cpu-baseline > $O/aa_bench.json 2> $O/aa_bench.err; echo "rc=$?"; tail -c 300 $O/aa_bench.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/aa_bench.json") if l.startswith("{")][-1])
print("ms_per_step", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), {k:round(v,2) for k,v in d["device_ms_per_step"].items()}, d["digest"]["matches_golden"])
for k,v in d.get("extras",{}).items(): print(" ", k, round(v.get("ms_per_step",0),2), v.get("digest",{}).get("matches_golden"), {a:round(b,2) for a,b in v.get("device_ms_per_step",{}).items()})
PY
