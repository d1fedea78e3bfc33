#!/bin/bash
# GP-coupled step: new GPU tests, then (optionally) the whole GPU suite and a bench line.  usage: tools/gpu_sur.sh <tag> [full]
tag=${1:-rS}; out=gpurun_out/$tag; mkdir -p $out
timeout 400 python -m pytest tests/test_surrogate_gpu.py -m gpu -q --timeout 200 > $out/sur.log 2>&1; echo "sur exit $?" | tee -a $out/sur.log
tail -60 $out/sur.log
if [ "$2" = "full" ]; then
  timeout 420 python -m pytest tests -m gpu -q --timeout 200 > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
  tail -8 $out/pytest.log
  timeout 120 python -c 'import __graft_entry__ as g; g.smoke()' > $out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 $out/smoke.log
  timeout 200 python bench.py --steps 200 --warmup 10 > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
  python -c "import json;d=json.load(open('$out/bench.json'));print(d['ms_per_step'],d['value'],d['roofline']['frac'],d['e2e']['value'])"
fi
