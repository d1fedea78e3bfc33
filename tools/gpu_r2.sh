#!/bin/bash
# One gpurun call of round 2: smoke, GPU parity tests, the bench line, ncu launch lists and --set full captures of the
# FV pass kernels (8192^2 journal) and of the GP kernels.   usage: tools/gpu_r2.sh <tag> [skip-tests]
tag=${1:-r2x}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.txt 2>&1
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $out/smoke.log 2>&1 || { echo SMOKE FAILED; tail -20 $out/smoke.log; exit 1; }
tail -3 $out/smoke.log
if [ -z "$2" ]; then
timeout 1500 python -m pytest tests -m gpu -q --timeout 400 --durations=10 > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -16 $out/pytest.log
fi
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
cat $out/bench.json | cut -c1-1500
tail -5 $out/bench.err
B="python bench.py --steps 3 --warmup 3 --no-cpu --no-extras --no-gp"
timeout 300 $B > $out/plain.log 2>&1 &&
GPF_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $out/launches.csv $B > $out/ncu_list.log 2>&1
timeout 300 $B > $out/plain2.log 2>&1 &&
GPF_NO_GRAPH=1 ncu --set full --clock-control none --import-source on -k regex:fv_pass -s 8 -c 4 -o $out/prof_fv $B > $out/ncu_full.log 2>&1
G="python tools/gp_profile.py"
timeout 300 $G > $out/gp_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gp_(var|mean)_kernel" -c 3 -o $out/prof_gp $G > $out/ncu_gp.log 2>&1
timeout 300 $G > $out/gp_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gp_gemm_kernel" -s 40 -c 1 -o $out/prof_gpsyrk $G > $out/ncu_gpsyrk.log 2>&1
timeout 300 $G > $out/gp_plain4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gp_gemm_kernel" -s 88 -c 4 -o $out/prof_gpinv $G > $out/ncu_gpinv.log 2>&1
timeout 300 $G > $out/gp_plain5.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gp_potrf_diag_kernel" -s 30 -c 1 -o $out/prof_gppotrf $G > $out/ncu_gppotrf.log 2>&1
timeout 300 $G > $out/gp_plain3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file $out/gp_launches.csv $G > $out/ncu_gplist.log 2>&1
ls -la $out
