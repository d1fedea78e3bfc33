#!/bin/bash
# Quick experiment: smoke (parity) + one short bench line.  usage: tools/gpu_quick.sh <tag>
tag=${1:-q}; out=gpurun_out/$tag; mkdir -p $out
timeout 120 python -c 'import __graft_entry__ as g; g.smoke()' > $out/smoke.log 2>&1 || { echo SMOKE FAILED; tail -20 $out/smoke.log; exit 1; }
tail -3 $out/smoke.log
timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
python - <<PY
import json
d=json.load(open("$out/bench.json"))
r=d["roofline"]
print("ms/step %.4f  value %.3f G/s  step_frac %.3f | pred %.4f ms (%.3f)  corr %.4f ms (%.3f) | e2e %.3f G/s" % (d["ms_per_step"], d["value"]/1e9, r["step_frac_of_hbm_roofline"], r["predictor"]["ms_per_launch"], r["predictor"]["frac"], r["ms_per_launch"], r["frac"], d["e2e"]["value"]/1e9))
PY
