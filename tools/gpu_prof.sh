#!/bin/bash
# ncu --set full of one predictor + one corrector launch.  usage: tools/gpu_prof.sh <tag>
tag=${1:-p}; out=gpurun_out/$tag; mkdir -p $out
timeout 120 python bench.py --steps 3 --warmup 3 --no-cpu > $out/plain.log 2>&1 &&
timeout 500 ncu --set full --clock-control none --import-source on -k regex:fv_pass -s 6 -c 2 -o $out/prof \
    python bench.py --steps 3 --warmup 3 --no-cpu > $out/ncu_full.log 2>&1
ls -la $out | tail -3
