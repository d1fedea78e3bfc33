#!/bin/bash
# run one test file on the GPU: tools/gpu_t.sh <tag> <pytest args...>
tag=$1; shift; out=gpurun_out/$tag; mkdir -p $out
timeout 600 python -m pytest "$@" -x -q --timeout 300 > $out/pytest.log 2>&1; echo "pytest exit $?"; tail -40 $out/pytest.log
