#!/bin/bash
# experiment: wave throttle window scan at 8192^2 + DRAM bytes under ncu.  usage: tools/gpu_exp.sh <tag>
tag=${1:-r2x}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_fv_gpu.py tests/test_slab_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x --timeout 400 > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -4 $out/pytest.log
B="python bench.py --steps 100 --warmup 10 --no-cpu --no-extras --no-gp"
for W in 0 1 2 3 5 8; do
  GPF_WAVE_WINDOW=$W timeout 300 $B > $out/bench_w$W.json 2> $out/bench_w$W.err
  python - <<PY
import json
d=json.load(open("$out/bench_w$W.json"))
r=d["roofline"]
print("W=$W ms/step %.4f pred %.4f corr %.4f exact %.4f clocks %s" % (d["ms_per_step"], r["predictor"]["ms_per_launch"], r["ms_per_launch"], d["other_arith"]["ms_per_step"], d["clocks"]))
PY
done
B2="python bench.py --steps 3 --warmup 3 --no-cpu --no-extras --no-gp"
for W in 0 3; do
GPF_WAVE_WINDOW=$W timeout 300 $B2 > $out/plain_w$W.log 2>&1 &&
GPF_WAVE_WINDOW=$W GPF_NO_GRAPH=1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:fv_pass -s 8 -c 4 --csv --log-file $out/dram_w$W.csv $B2 > $out/ncu_w$W.log 2>&1
grep fv_pass $out/dram_w$W.csv | awk -F'","' '{print $5, $(NF-2), $(NF-1), $NF}' | cut -c1-200
done
