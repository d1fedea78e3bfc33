#!/bin/bash
tag=$1
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_gp_gpu.py tests/test_surrogate_gpu.py tests/test_fullsize_gpu.py -m gpu -q --timeout 300 -x > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -8 $out/pytest.log
timeout 600 python bench_gp.py --no-cpu --coupled-n 0 > $out/bench_gp.json 2> $out/bench_gp.err; echo "bench_gp exit $?"
python - <<PY
import json
d=json.load(open("$out/bench_gp.json"))
for k,m in d["models"].items():
    print(k, "factor %.2f ms nlml %.2f ms (%.1f TF) mean %.3f ms var %s" % (m["factor_ms"], m["nlml_grad_ms"], m["nlml_grad_tflops"], m["mean"]["ms"], {a:b for a,b in m["mean_var"].items() if a in ("ms","N","var_kernel_tflops","frac_of_dgemm_peak")}))
PY
G="python tools/gp_profile.py"
timeout 300 $G > $out/gp_plain3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file $out/gp_launches.csv $G > $out/ncu_gplist.log 2>&1
