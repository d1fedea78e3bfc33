#!/bin/bash
tag=${1:-gp}; out=gpurun_out/$tag; mkdir -p $out
timeout 600 python -m pytest tests/test_gp_gpu.py -x -q --timeout 300 > $out/pytest_gp.log 2>&1; echo "pytest exit $?"; tail -30 $out/pytest_gp.log
