#!/bin/bash
# pipelined host-buffer step: its test, then the bench line (e2e).  usage: tools/gpu_host.sh <tag>
tag=${1:-rH}; out=gpurun_out/$tag; mkdir -p $out
timeout 300 python -m pytest tests/test_fv_gpu.py -m gpu -q --timeout 200 -k "pipelined or host_update" > $out/host.log 2>&1; echo "host tests exit $?" | tee -a $out/host.log
tail -30 $out/host.log
timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu > $out/bench.json 2> $out/bench.err; echo "bench exit $?"; tail -3 $out/bench.err
python -c "import json;d=json.load(open('$out/bench.json'));print(d['ms_per_step'],d['value'],d['roofline']['frac'],d['e2e'])"
GPF_HOST_CHUNKS=16 timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu 2>/dev/null | python -c "import json,sys;d=json.loads(sys.stdin.read());print('chunks16',d['e2e'])"
GPF_HOST_CHUNKS=4 timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu 2>/dev/null | python -c "import json,sys;d=json.loads(sys.stdin.read());print('chunks4',d['e2e'])"
GPF_HOST_CHUNKS=1 timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu 2>/dev/null | python -c "import json,sys;d=json.loads(sys.stdin.read());print('plain',d['e2e'])"
