#!/bin/bash
tag=$1; N=$2
out=gpurun_out/$tag
mkdir -p $out
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -6 $out/pytest.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
run() { # name, env, args
  env $2 timeout 600 $T bench.py --gpus $N --steps 100 --warmup 10 $3 > $out/$1.json 2> $out/$1.err; echo "$1 exit $?"
  python - <<PY
import json
try:
    d=json.load(open("$out/$1.json"))
    print("$1", "ms/step %.4f value %.4g G/s launches %d parity %s clocks %s" % (d["ms_per_step"], d["value"]/1e9, d["gpu_launches"], (d.get("slab_parity") or {}).get("bit_exact"), d["clocks"]))
except Exception as e:
    print("$1 failed", e); print(open("$out/$1.err").read()[-2000:])
PY
}
run fused4096 "GPF_X=1" "--cells 4096"
run fused8192 "GPF_X=1" "--no-parity"
python bench_gp.py --no-cpu --coupled-n 1024 --coupled-steps 3 > $out/bench_gp.json 2> $out/bench_gp.err; echo "bench_gp exit $?"
python - <<PY
import json
d=json.load(open("$out/bench_gp.json"))
for k,m in d["models"].items():
    print(k, "factor %.2f ms nlml %.2f ms mean %.3f ms var %s" % (m["factor_ms"], m["nlml_grad_ms"], m["mean"]["ms"], {a:b for a,b in m["mean_var"].items() if a in ("ms","N","var_kernel_tflops","frac_of_dgemm_peak")}))
print("coupled", d["coupled_step"]["ms_per_step"])
PY
