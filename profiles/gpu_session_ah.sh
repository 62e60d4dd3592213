#!/bin/bash
# Round-2 GPU session AH (1 GPU): final code — suite, smoke, headline line without extras.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1500 python -m pytest tests -m gpu -q > $O/ah_pytest.log 2>&1; echo "rc=$?" >> $O/ah_pytest.log; tail -4 $O/ah_pytest.log
echo "== smoke"; python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
echo "== bench c2"; timeout 600 python bench.py --extras none > $O/ah_bench.json 2> $O/ah_bench.err; echo "rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/ah_bench.json") if l.startswith("{")][-1])
print("ms_per_step", round(d["ms_per_step"],2), "value", round(d["value"]/1e6,2), "e2e", round(d["e2e"]["value"]/1e6,2), round(d["e2e"]["ms_per_step"],2), d["digest"]["matches_golden"], "frac", round(d["roofline"]["frac"],3))
PY
