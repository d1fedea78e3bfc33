#!/bin/bash
# GPU tests with the log kept: tools/gpu_t.sh <tag> [pytest args...]
tag=$1; shift
out=gpurun_out/$tag
mkdir -p $out
timeout 1200 python -m pytest "$@" -m gpu -q --timeout 120 --durations=10 > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -25 $out/pytest.log
