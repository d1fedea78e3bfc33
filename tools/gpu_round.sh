#!/bin/bash
# One gpurun call: GPU parity tests, bench line, ncu launch list + full capture of the pass kernels.
# usage: tools/gpu_round.sh <tag>
tag=${1:-rX}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.txt 2>&1
timeout 120 python -c 'import __graft_entry__ as g; g.smoke()' > $out/smoke.log 2>&1 || { echo SMOKE FAILED; tail -20 $out/smoke.log; exit 1; }
tail -3 $out/smoke.log
timeout 600 python -m pytest tests -m gpu -x -q --timeout 200 > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -3 $out/pytest.log
grep -q "pytest exit 0" $out/pytest.log || { tail -40 $out/pytest.log; exit 1; }
timeout 300 python bench.py --steps 200 --warmup 10 > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
cat $out/bench.json
timeout 120 python bench.py --steps 3 --warmup 3 --no-cpu > $out/plain.log 2>&1 &&
GPF_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $out/launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu > $out/ncu_list.log 2>&1
timeout 120 python bench.py --steps 3 --warmup 3 --no-cpu > $out/plain2.log 2>&1 &&
GPF_NO_GRAPH=1 ncu --set full --clock-control none --import-source on -k regex:fv_pass -s 8 -c 4 -o $out/prof \
    python bench.py --steps 3 --warmup 3 --no-cpu > $out/ncu_full.log 2>&1
ls -la $out
