#!/bin/bash
# Round-2 GPU session AC (1 GPU): final code — suite, default bench (headline + extras at full size), reference arm, smoke, launch list of one pass.
set -u
O=gpurun_out
mkdir -p $O
echo "== pytest"; timeout 1500 python -m pytest tests -m gpu -q > $O/ac_pytest.log 2>&1; echo "rc=$?" >> $O/ac_pytest.log; tail -4 $O/ac_pytest.log
echo "== bench default"; timeout 1200 python bench.py > $O/ac_bench.json 2> $O/ac_bench.err; echo "rc=$?"; tail -c 400 $O/ac_bench.err
echo "== reference arm"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/ac_bench_ref.json 2> $O/ac_bench_ref.err; echo "rc=$?"
echo "== smoke"; python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
echo "== ncu launch list (one step incl. warm-up pass)"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $O/ac_launches.csv python bench.py --steps 1 --warmup 1 --extras none --no-cpu-baseline --skip-e2e --no-digest > $O/ac_ncu_list.log 2>&1; echo "rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/ac_bench.json") if l.startswith("{")][-1])
print("ms_per_step", round(d["ms_per_step"],2), "value", round(d["value"]/1e6,2), "e2e", round(d["e2e"]["value"]/1e6,2), round(d["e2e"]["ms_per_step"],2), {k:round(v,2) for k,v in d["device_ms_per_step"].items()}, d["digest"]["matches_golden"], "frac", round(d["roofline"]["frac"],3), "cpu", d["cpu_baseline"]["value"])
for k,v in d.get("extras",{}).items(): print(" ", k, round(v.get("ms_per_step",0),2), v.get("digest",{}).get("matches_golden"), {a:round(b,1) for a,b in v.get("device_ms_per_step",{}).items()}, v.get("roofline",{}).get("frac"))
PY
