#!/bin/bash
# ncu --set full of the exact-contract pass kernels. usage: tools/gpu_prof_exact.sh <tag>
tag=${1:-px}; out=gpurun_out/$tag; mkdir -p $out
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu --arith exact > $out/plain.log 2>&1 || { tail -5 $out/plain.log; exit 1; }
GPF_NO_GRAPH=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:fv_pass -s 8 -c 4 -o $out/prof \
    python bench.py --steps 3 --warmup 3 --no-cpu --arith exact > $out/ncu_full.log 2>&1
ls -la $out
