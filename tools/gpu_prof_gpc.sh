#!/bin/bash
# ncu evidence for the GP-coupled step: launch list of one small run, then --set full of the mean kernels
# (D = 1 and D = 3 variants) and of the surrogate pass kernels.  usage: tools/gpu_prof_gpc.sh <tag>
tag=${1:-pg}; out=gpurun_out/$tag; mkdir -p $out
ARGS="--coupled-only --no-cpu --coupled-n 1024 --M 4096 --coupled-steps 2"
timeout 200 python bench_gp.py $ARGS > $out/plain.json 2> $out/plain.err || { echo PLAIN FAILED; tail -5 $out/plain.err; exit 1; }
cat $out/plain.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $out/launches.csv \
    python bench_gp.py $ARGS > $out/ncu_list.log 2>&1; echo "list exit $?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'gp_mean_kernel|fv_pass_kernel' -s 20 -c 9 -o $out/prof \
    python bench_gp.py $ARGS > $out/ncu_full.log 2>&1; echo "full exit $?"
ncu -i $out/prof.ncu-rep --page raw --csv > $out/prof_raw.csv 2> /dev/null; echo "raw exit $?"
ls -la $out | tail -8
