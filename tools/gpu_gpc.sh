#!/bin/bash
# GP kernels + coupled step: GPU tests of both, then bench_gp.py (with the coupled step).  usage: tools/gpu_gpc.sh <tag>
tag=${1:-rG}; out=gpurun_out/$tag; mkdir -p $out
timeout 400 python -m pytest tests/test_gp_gpu.py tests/test_surrogate_gpu.py -m gpu -q --timeout 200 > $out/gp.log 2>&1; echo "gp tests exit $?" | tee -a $out/gp.log
tail -25 $out/gp.log
timeout 500 python bench_gp.py > $out/bench_gp.json 2> $out/bench_gp.err; echo "bench_gp exit $?"
tail -5 $out/bench_gp.err
python - <<PY
import json
d=json.load(open("$out/bench_gp.json"))
for k,m in d["models"].items(): print(k, "mean ms %.2f (%.1f M pred/s)  mean_var ms %.1f  factor ms %.1f" % (m["mean"]["ms"], m["mean"]["value"]/1e6, m["mean_var"]["ms"], m["factor_ms"]))
print(json.dumps(d.get("coupled_step"), indent=1))
PY
