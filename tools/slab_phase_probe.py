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
names = ["pass0", "pack1", "xchg1", "unpack1", "pass1", "pack2", "xchg2", "unpack2", "record", "gather", "finish"]
tot = {k: 0.0 for k in names}
N = 50
evs = []
for it in range(N):
    e = [torch.cuda.Event(enable_timing=True) for _ in range(len(names) + 1)]
    e[0].record()
    s._pass(0); e[1].record()
    s._pack(1); e[2].record()
    s.ex.exchange(s.send, s.recv); e[3].record()
    s._unpack(1); e[4].record()
    s._pass(1); e[5].record()
    s._pack(2); e[6].record()
    s.ex.exchange(s.send, s.recv); e[7].record()
    s._unpack(2); e[8].record()
    s._record(0); e[9].record()
    s.ex.gather(s.rec, s.recs); e[10].record()
    s._finish(_capi.FIN_STEP); e[11].record()
    evs.append(e)
torch.cuda.synchronize()
for e in evs:
    for k, nm in enumerate(names):
        tot[nm] += e[k].elapsed_time(e[k + 1])
if rank == 0:
    print("per-step phase times (us), rank 0, %d steps:" % N)
    for nm in names:
        print("  %-8s %8.1f" % (nm, 1e3 * tot[nm] / N))
    print("  total    %8.1f" % (1e3 * sum(tot.values()) / N))
dist.barrier(); dist.destroy_process_group()
