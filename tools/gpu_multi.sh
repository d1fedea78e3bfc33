#!/bin/bash
# multi-GPU round: NCCL / IPC tests, bench at N ranks (headline + small config, fused and with the exchange kernels).
# usage: tools/gpu_multi.sh <tag> <N>
tag=$1; N=$2
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=index,name --format=csv > $out/smi.txt
timeout 900 python -m pytest tests/test_slab_gpu.py tests/test_surrogate_gpu.py tests/test_slip_closures.py -m gpu -q --timeout 600 -k "nccl or ipc or slip or device" > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -5 $out/pytest.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
run() { # name, env, args
  env $2 timeout 900 $T bench.py --gpus $N --steps 200 --warmup 20 $3 > $out/$1.json 2> $out/$1.err; echo "$1 exit $?"
  python - <<PY
import json
try:
    d=json.load(open("$out/$1.json"))
    print("$1", "ms/step %.4f value %.4g G/s launches %d parity %s clocks %s notes %s" % (d["ms_per_step"], d["value"]/1e9, d["gpu_launches"], (d.get("slab_parity") or {}).get("bit_exact"), d["clocks"], d["notes"]))
    if d.get("gp"): print("   gp coupled", d["gp"]["coupled_step"]["ms_per_step"], d["gp"]["coupled_step"]["cell_updates_per_s"])
except Exception as e:
    print("$1 failed", e); print(open("$out/$1.err").read()[-2000:])
PY
}
run n${N}_journal8192 "GPF_X=1" ""
run n${N}_journal4096_fused "GPF_X=1" "--cells 4096 --no-parity --no-gp"
run n${N}_journal4096_kernels "GPF_SLAB_FUSED=0" "--cells 4096 --no-parity --no-gp"
