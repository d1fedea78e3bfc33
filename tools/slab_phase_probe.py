"""Per-phase CUDA-event timing of the slab step (non-overlapped):
torchrun --nproc-per-node 2 tools/slab_phase_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import bench
from gapflow_b200 import _capi
from gapflow_b200.dist import SlabSolver, slab_rows
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
n = 4096
d = bench.slider_inputs(n * world, n)
start, nloc = slab_rows(d['grid']['Nx'], world)[rank]
q0, topo = bench.initial_fields(d, rows=(start, start + nloc + 2))
s = SlabSolver(d['grid'], d['geometry'], d['properties'], d['numerics'], topo=topo, q=q0, rank=rank, world=world,
               arith='fast', overlap=False)
s.fill_ghosts(); s.pre_run(False)
s.step(20)
torch.cuda.synchronize(); dist.barrier()
import ctypes as C
peer = s._comm is not None
lo, hi = s._neighbours()
st = s._s
if peer:
    names = ["pass0", "send1", "recv1", "pass1", "send2", "recv2", "rec_send", "fin_recv"]
    calls = [lambda: s._pass(0),
             lambda: _capi.check(s.lib.gpf_fv_halo_send(s.h, s._comm, 1, lo, hi, st())),
             lambda: _capi.check(s.lib.gpf_fv_halo_recv(s.h, s._comm, 1, st())),
             lambda: s._pass(1),
             lambda: _capi.check(s.lib.gpf_fv_halo_send(s.h, s._comm, 2, lo, hi, st())),
             lambda: _capi.check(s.lib.gpf_fv_halo_recv(s.h, s._comm, 2, st())),
             lambda: _capi.check(s.lib.gpf_fv_record_send(s.h, s._comm, 0, st())),
             lambda: _capi.check(s.lib.gpf_fv_finish_recv(s.h, s._comm, _capi.FIN_STEP, C.c_void_p(s.recs.data_ptr()), st()))]
else:
    names = ["pass0", "pack1", "xchg1", "unpack1", "pass1", "pack2", "xchg2", "unpack2", "record", "gather", "finish"]
    calls = [lambda: s._pass(0), lambda: s._pack(1), lambda: s.ex.exchange(s.send, s.recv), lambda: s._unpack(1),
             lambda: s._pass(1), lambda: s._pack(2), lambda: s.ex.exchange(s.send, s.recv), lambda: s._unpack(2),
             lambda: s._record(0), lambda: s.ex.gather(s.rec, s.recs), lambda: s._finish(_capi.FIN_STEP)]
tot = {k: 0.0 for k in names}
N = 50
evs = []
for it in range(N):
    e = [torch.cuda.Event(enable_timing=True) for _ in range(len(names) + 1)]
    e[0].record()
    for k, f in enumerate(calls):
        f()
        e[k + 1].record()
    evs.append(e)
torch.cuda.synchronize()
for e in evs:
    for k, nm in enumerate(names):
        tot[nm] += e[k].elapsed_time(e[k + 1])
if rank == 0:
    print("per-step phase times (us), rank 0, %d steps, %s:" % (N, "peer memory" if peer else "NCCL"))
    for nm in names:
        print("  %-8s %8.1f" % (nm, 1e3 * tot[nm] / N))
    print("  total    %8.1f" % (1e3 * sum(tot.values()) / N))
dist.barrier(); dist.destroy_process_group()
