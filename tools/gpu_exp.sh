#!/bin/bash
# Timing-only experiment (results may be wrong on purpose): one short bench line.  usage: tools/gpu_exp.sh <tag>
tag=${1:-x}; out=gpurun_out/$tag; mkdir -p $out
timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu > $out/bench.json 2> $out/bench.err; echo "bench exit $?"; tail -3 $out/bench.err
python - <<PY
import json
d=json.load(open("$out/bench.json"))
r=d["roofline"]
print("ms/step %.4f  value %.3f G/s  step_frac %.3f | pred %.4f ms (%.3f)  corr %.4f ms (%.3f)" % (d["ms_per_step"], d["value"]/1e9, r["step_frac_of_hbm_roofline"], r["predictor"]["ms_per_launch"], r["predictor"]["frac"], r["ms_per_launch"], r["frac"]))
PY
