#!/bin/bash
# Last-budget GPU call: the newest tests first, then smoke, the full GPU suite, one bench line.
tag=${1:-rZ}; out=gpurun_out/$tag; mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.txt 2>&1
timeout 200 python -m pytest tests/test_fv_gpu.py -m gpu -q -k "viscosity" --timeout 150 > $out/visc.log 2>&1; echo "visc exit $?" | tee -a $out/visc.log
tail -15 $out/visc.log
timeout 120 python -c 'import __graft_entry__ as g; g.smoke()' > $out/smoke.log 2>&1; echo "smoke exit $?" | tee -a $out/smoke.log
tail -4 $out/smoke.log
timeout 200 python bench.py --steps 200 --warmup 10 > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
cat $out/bench.json
timeout 420 python -m pytest tests -m gpu -q --timeout 200 > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -25 $out/pytest.log
