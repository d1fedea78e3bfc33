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
echo "profiles: tools/gpu_prof.sh"
