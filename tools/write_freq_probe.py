"""Problem.run() with a small write_freq: the run loop's frames staged through a side stream (async_write, default)
against synchronous writes and against no output.   python tools/write_freq_probe.py [n] [steps] [write_freq]"""
import os
import shutil
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from gapflow_b200.problem import Problem

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
wf = int(sys.argv[3]) if len(sys.argv) > 3 else 10
tmp = tempfile.mkdtemp(prefix="gpf_wf_")
res = {}
for mode in ("silent", "sync", "async", "sync", "async"):
    d = bench.journal_inputs(n, n)
    d['options'].update(output=os.path.join(tmp, mode + str(len(res))), use_tstamp=False, write_freq=wf, silent=mode == "silent",
                        async_write=mode == "async")
    d['numerics'].update(max_it=steps, tol=1e-30)
    p = Problem._from_dict(d)
    p._pre_run()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sys.stdout = open(os.devnull, "w")
    try:
        p.run()
    finally:
        sys.stdout = sys.__stdout__
    torch.cuda.synchronize()
    res.setdefault(mode, []).append(time.perf_counter() - t0)
    shutil.rmtree(d['options']['output'], ignore_errors=True)
shutil.rmtree(tmp, ignore_errors=True)
frame_mb = 28 * (n + 2) ** 2 * 8 / 1e6
print(f"journal {n}x{n}, {steps} steps, write_freq {wf} ({steps // wf + 1} frames of {frame_mb:.0f} MB): "
      + ", ".join(f"{k} {min(v):.3f} s" for k, v in res.items()))
