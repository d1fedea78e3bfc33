#!/bin/bash
# multi-GPU round: NCCL / IPC tests, bench at N ranks fused and with the exchange kernels.  usage: tools/gpu_multi.sh <tag> <N>
tag=$1; N=$2
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=index,name --format=csv > $out/smi.txt
timeout 900 python -m pytest tests/test_slab_gpu.py tests/test_surrogate_gpu.py -m gpu -q --timeout 600 -k "nccl or ipc" > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -5 $out/pytest.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
run() { # name, env, args
  env $2 timeout 600 $T bench.py --gpus $N --steps 100 --warmup 10 $3 > $out/$1.json 2> $out/$1.err; echo "$1 exit $?"
  python - <<PY
import json
try:
    d=json.load(open("$out/$1.json"))
    print("$1", "ms/step %.4f value %.4g G/s launches %d parity %s clocks %s" % (d["ms_per_step"], d["value"]/1e9, d["gpu_launches"], d.get("slab_parity"), d["clocks"]))
except Exception as e:
    print("$1 failed", e); print(open("$out/$1.err").read()[-2000:])
PY
}
run fused8192 "GPF_X=1" ""
run unfused8192 "GPF_SLAB_FUSED=0" "--no-parity"
run fused4096 "GPF_X=1" "--cells 4096 --no-parity"
run unfused4096 "GPF_SLAB_FUSED=0" "--cells 4096 --no-parity"
