"""Dry run of bench.py's N > 1 control flow on the CPU (helper of tests/test_host_cpu.py, not a test module).

What it checks is HOST logic only: the order of the collectives on every rank, the assembly of the JSON line, the
watchdogs.  The solver classes are replaced by stand-ins that keep their field unchanged (so slab_parity's comparison
of the slabs with the single domain also checks that bench.initial_fields gives every decomposition the same global
field), "cuda" tensors are created on the host and the process group is gloo.  No number it prints means anything.

    RANK=r WORLD_SIZE=2 LOCAL_RANK=r MASTER_ADDR=127.0.0.1 MASTER_PORT=p python tests/bench_dryrun.py <bench args>
    DRYRUN_HANG=parity: rank 1 never returns from the first step of slab_parity (a peer that hangs in collective code)
    DRYRUN_HANG=main:   rank 1 never returns from the warm-up steps of the measurement
"""
import os
import sys
import time
import types

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import gapflow_b200.dist as gdist  # noqa: E402
import gapflow_b200.fv as gfv  # noqa: E402


def _on_host(fn):
    def wrapped(*a, **k):
        if str(k.get('device', '')).startswith('cuda'):
            k['device'] = 'cpu'
        k.pop('pin_memory', None)
        return fn(*a, **k)
    return wrapped


for _name in ('tensor', 'empty', 'zeros', 'as_tensor'):
    setattr(torch, _name, _on_host(getattr(torch, _name)))


class _Event:
    def __init__(self, *a, **k):
        self.t = None

    def record(self, *a):
        self.t = time.perf_counter()

    def elapsed_time(self, other):
        return (other.t - self.t) * 1e3 + 1.0


torch.cuda.is_available = lambda: True
torch.cuda.set_device = lambda *a, **k: None
torch.cuda.synchronize = lambda *a, **k: None
torch.cuda.empty_cache = lambda: None
torch.cuda.Event = _Event
_init = dist.init_process_group
dist.init_process_group = lambda backend=None, device_id=None, **k: _init("gloo", **k)

IN_PARITY = {"on": False}
_parity = bench.slab_parity


def _slab_parity(*a, **k):
    IN_PARITY["on"] = True
    return _parity(*a, n=int(os.environ.get("DRYRUN_PARITY_N", "64")), **k)


bench.slab_parity = _slab_parity


class _Solver:
    """Keeps q as given; counts steps; fixed dt and kinetic energy."""

    def __init__(self, grid, geo, prop, numerics, topo=None, q=None, rank=0, world=1, arith='exact', **kw):
        self.q = None if q is None else np.array(q)
        self.rank, self.world, self.steps, self.closed = rank, world, 0, False
        self.fv = self

    def set_q(self, q):
        self.q = np.array(q)

    def fill_ghosts(self):
        pass

    def pre_run(self, reset):
        pass

    def step(self, n=1, until_converged=False):
        hang = os.environ.get("DRYRUN_HANG", "")
        if self.world > 1 and self.rank == 1 and hang == ("parity" if IN_PARITY["on"] else "main"):
            time.sleep(3600)
        self.steps += n

    def scalars(self):
        return types.SimpleNamespace(step=self.steps, stop=False, dt=1e-9, ekin=2.5, residual=1e-3)

    def comm_error(self):
        return False

    def set_timing(self, on):
        pass

    def timing(self):
        return 1.0, 1.0, 0

    def get_q(self, out=None):
        if out is None:
            return self.q.copy()
        out[...] = torch.from_numpy(self.q) if isinstance(out, torch.Tensor) else self.q
        return out

    def update_host(self, qh, n):
        self.steps += n

    def close(self):
        assert not self.closed
        self.closed = True
        if self.world > 1:
            dist.barrier()


gdist.SlabSolver = _Solver
gfv.FVSolver = _Solver

if __name__ == "__main__":
    bench.main()
