#!/bin/bash
# tests + default bench line + GP numbers.  usage: tools/gpu_round2.sh <tag>
tag=$1
out=gpurun_out/$tag
mkdir -p $out
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -6 $out/pytest.log
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
python - <<PY
import json
d=json.load(open("$out/bench.json"))
r=d["roofline"]
print("ms/step %.4f value %.4g frac %.4f dram_frac %s step_frac %.4f step_dram_frac %s pred %.4f clocks %s" % (d["ms_per_step"], d["value"], r["frac"], r["dram_frac"], r["step_frac_of_hbm_roofline"], r["step_dram_frac"], r["predictor"]["frac"], d["clocks"]))
print("e2e %.4g other %s" % (d["e2e"]["value"], d["other_arith"]))
print("slider", d["slider4096"]); print("variants", d["variants"]); print("notes", d["notes"])
g=d["gp"]
for k,m in g["models"].items():
    print(k, "factor %.2f ms nlml %.2f ms mean %.3f ms var %s" % (m["factor_ms"], m["nlml_grad_ms"], m["mean"]["ms"], {a:b for a,b in m["mean_var"].items() if a in ("ms","N","var_kernel_tflops","frac_of_dgemm_peak")}))
print("coupled", g["coupled_step"]["ms_per_step"], "parity", g.get("parity_vs_cpu_port"))
PY
tail -3 $out/bench.err
