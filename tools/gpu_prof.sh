#!/bin/bash
# ncu captures of the round's kernels, reduced on the box to raw-page CSVs (the .ncu-rep files exceed what gpurun
# carries back).   usage: tools/gpu_prof.sh <tag>
tag=${1:-r2x}
out=gpurun_out/$tag
mkdir -p $out
cap() { # name, command..., then ncu args after --
  name=$1; shift
  ncu --set full --clock-control none --import-source on -o $out/$name "$@" > $out/ncu_$name.log 2>&1
  ncu -i $out/$name.ncu-rep --page raw --csv > $out/$name.csv 2>/dev/null
  rm -f $out/$name.ncu-rep
}
B="python bench.py --steps 3 --warmup 3 --no-cpu --no-extras --no-gp"
timeout 300 $B > $out/plain.log 2>&1 &&
GPF_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $out/launches.csv $B > $out/ncu_list.log 2>&1
timeout 300 $B > $out/plain2.log 2>&1 && GPF_NO_GRAPH=1 cap prof_fv -k regex:fv_pass -s 8 -c 4 $B
G="python tools/gp_profile.py"
timeout 300 $G > $out/gp_plain.log 2>&1 && cap prof_gp -k regex:"gp_(var|mean)_kernel" -c 3 $G
timeout 300 $G > $out/gp_plain2.log 2>&1 && cap prof_gpsyrk -k regex:"gp_gemm_kernel" -s 40 -c 1 $G
timeout 300 $G > $out/gp_plain4.log 2>&1 && cap prof_gpinv -k regex:"gp_gemm_kernel" -s 88 -c 4 $G
timeout 300 $G > $out/gp_plain5.log 2>&1 && cap prof_gppotrf -k regex:"gp_potrf_diag_kernel" -s 30 -c 1 $G
timeout 300 $G > $out/gp_plain3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file $out/gp_launches.csv $G > $out/ncu_gplist.log 2>&1
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
du -sh $out
