#!/bin/bash
# Multi-GPU check: slab tests (in-process + NCCL) and the weak-scaling bench at N = 1, 2[, 4, 8].
tag=${1:-rXm}; shift
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi -L > $out/smi.txt
timeout 600 python -m pytest tests/test_slab_gpu.py -x -q --timeout 300 > $out/pytest_slab.log 2>&1; echo "pytest exit $?"; tail -15 $out/pytest_slab.log
for n in "$@"; do
  if [ "$n" = 1 ]; then
    timeout 300 python bench.py --gpus 1 --steps 100 --warmup 10 --no-cpu > $out/bench_n1.json 2> $out/bench_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus $n --steps 100 --warmup 10 > $out/bench_n$n.json 2> $out/bench_n$n.err
  fi
  echo "bench n=$n exit $?"; tail -c 1500 $out/bench_n$n.json; tail -5 $out/bench_n$n.err
done
