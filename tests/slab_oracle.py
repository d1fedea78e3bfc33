"""CPU model of the slab decomposition: the NumPy oracle run on x-slabs, with halo rows and
loop scalars exchanged through gapflow_b200.dist.HaloExchanger on CPU tensors (gloo).  Test
infrastructure: pins the partition / neighbour / message-order logic of dist.py without a GPU."""
import numpy as np
import torch

from gapflow_b200.dist import HaloExchanger, local_grid, slab_rows, x_periodic
from oracle.fv import FVOracle
from oracle import closures as C


class SlabOracle(FVOracle):
    def __init__(self, d, q_global, topo_global, extra_global, rank, world, group=None):
        g, start = local_grid(d['grid'], rank, world)
        n = g['Nx']
        sl = slice(start, start + n + 2)
        self.ex = HaloExchanger(rank, world, x_periodic(d['grid']), group)
        self.world = world
        # halo rows belong to the neighbour, except the ring's two ends: those are the global ghost rows
        self.own = slice(1 if (self.ex.lo is not None and rank != 0) else 0,
                         -1 if (self.ex.hi is not None and rank != world - 1) else None)
        self._ready = False
        super().__init__(g, d['geometry'], d['properties'], d['numerics'],
                         extra=None if extra_global is None else np.asarray(extra_global)[..., sl, :],
                         topo=topo_global[:, sl])
        self._ready = True
        self.kinetic_energy_old = self.kinetic_energy     # of the _initialize state (problem.py:670)
        self.q[...] = q_global[:, sl]
        if d['numerics']['adaptive']:
            self.dt = d['numerics']['CFL'] * self.dt_crit

    # -- scalars folded over ranks in rank order (dist.py: one record per rank, all-gathered) ----
    def _gather(self, vals):
        rec = torch.tensor(vals, dtype=torch.float64)
        out = torch.zeros((self.world, len(vals)), dtype=torch.float64)
        self.ex.gather(rec, out)
        return out.numpy()

    @property
    def kinetic_energy(self):
        q = self.q[:, self.own]
        loc = np.sum((q[1]**2 + q[2]**2) / q[0] / 2.)
        if not self._ready:
            return loc
        tot = 0.0
        for v in self._gather([loc])[:, 0]:
            tot += v
        return tot

    @property
    def v_max(self):
        q = self.q[:, self.own]
        return self._gather([np.sqrt((q[1]**2 + q[2]**2) / q[0]).max()])[:, 0].max()

    @property
    def v_sound(self):
        return self._gather([C.eos_sound_velocity(self.q[0, self.own], self.prop).max()])[:, 0].max()

    @property
    def q_is_valid(self):
        q = self.q[:, self.own]
        bad = float(np.any(np.isnan(q)) or np.any(q[0] < 0.))
        return self._gather([bad])[:, 0].max() == 0.

    # -- ghost cells: physical sides locally (x first, then y), slab-internal rows by exchange ----
    def fill_ghosts(self):
        g, q = self.grid, self.q
        if self.ex.lo is None:
            if all(g['bc_xE_P']):
                q[:, 0, :] = q[:, -2, :].copy()
            else:
                q[g['bc_xE_D'], :1, :] = self.ghost_values('D', 0, -1)
                q[g['bc_xE_N'], :1, :] = self.ghost_values('N', 0, -1)
        if self.ex.hi is None:
            if all(g['bc_xW_P']):
                q[:, -1, :] = q[:, 1, :].copy()
            else:
                q[g['bc_xW_D'], -1:, :] = self.ghost_values('D', 0, 1)
                q[g['bc_xW_N'], -1:, :] = self.ghost_values('N', 0, 1)
        if all(g['bc_yS_P']):
            q[:, :, 0] = q[:, :, -2].copy()
        else:
            q[g['bc_yS_D'], :, :1] = self.ghost_values('D', 1, -1)
            q[g['bc_yS_N'], :, :1] = self.ghost_values('N', 1, -1)
        if all(g['bc_yN_P']):
            q[:, :, -1] = q[:, :, 1].copy()
        else:
            q[g['bc_yN_D'], :, -1:] = self.ghost_values('D', 1, 1)
            q[g['bc_yN_N'], :, -1:] = self.ghost_values('N', 1, 1)
        send = torch.from_numpy(np.ascontiguousarray(np.stack([q[:, 1, :], q[:, -2, :]])))
        recv = torch.zeros_like(send)
        self.ex.exchange(send, recv)
        if self.ex.lo is not None:
            q[:, 0, :] = recv[0].numpy()
        if self.ex.hi is not None:
            q[:, -1, :] = recv[1].numpy()

    def gather_q(self, Nx_global):
        import torch.distributed as dist
        parts = [None] * self.world
        dist.all_gather_object(parts, self.q.copy(), group=self.ex.group)
        out = np.empty((3, Nx_global + 2, self.q.shape[2]))
        for r, (start, n) in enumerate(slab_rows(Nx_global, self.world)):
            lo = 0 if r == 0 else 1
            hi = n + 2 if r == self.world - 1 else n + 1
            out[:, start + lo:start + hi] = parts[r][:, lo:hi]
        return out
