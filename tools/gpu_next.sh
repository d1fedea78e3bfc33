#!/bin/bash
# First multi-GPU call of a next round: what round 2 could not measure after its GPU budget was gone.
# usage (under gpurun --gpus N): tools/gpu_next.sh <tag> <N>
tag=$1; N=${2:-8}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=index,name --format=csv > $out/smi.txt
# 1. the multi-process tests (one GPU per rank), then the whole suite once
timeout 900 python -m pytest tests/test_slab_gpu.py tests/test_surrogate_gpu.py tests/test_problem_gpu.py -m gpu -q --timeout 600 \
    -k "nccl or ipc or slabs" > $out/pytest_multi.log 2>&1; echo "pytest multi exit $?" | tee -a $out/pytest_multi.log
tail -3 $out/pytest_multi.log
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29533"
run() { # name, ranks, env, args      every run bounded three times: bench's own deadlines, then timeout
  env $3 timeout 700 $T --nproc-per-node $2 bench.py --gpus $2 $4 > $out/$1.json 2> $out/$1.err; echo "$1 exit $?"
  python - <<PY
import json
try:
    d = json.load(open("$out/$1.json"))
    print("$1", d.get("ms_per_step"), d.get("value"), "launches", d.get("gpu_launches"), "parity", d.get("slab_parity"), "error", d.get("error"), "clocks", d.get("clocks"))
except Exception as e:
    print("$1: no line:", e); print(open("$out/$1.err").read()[-1500:])
PY
}
# 2. the headline at every N of the driver's scaling run, measurement only, then with slab_parity
for n in 2 4 8; do [ $n -le $N ] && run n${n}_8192_noparity $n "GPF_X=1" "--no-parity"; done
for n in 2 4 8; do [ $n -le $N ] && run n${n}_8192 $n "GPF_X=1" ""; done
# 3. fused exchange against the exchange kernels at the size where the exchange shows
for n in 2 4 8; do [ $n -le $N ] && run n${n}_4096_fused $n "GPF_X=1" "--cells 4096 --no-parity" && run n${n}_4096_kernels $n "GPF_SLAB_FUSED=0" "--cells 4096 --no-parity"; done
# 4. the GP-coupled step on slabs: validated at N = 2 only so far
for n in 2 4 8; do [ $n -le $N ] && run n${n}_gp_slabs $n "GPF_X=1" "--cells 4096 --no-parity --gp-slabs"; done
du -sh $out
