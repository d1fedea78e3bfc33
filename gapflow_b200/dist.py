"""Slab decomposition of the FV step along x: one process per GPU, halo rows and loop scalars
exchanged with ``torch.distributed`` (NCCL over NVLink on the GPUs, gloo in the CPU tests).

The reference is single-process (GaPFlow/problem.py:122-124; its ``_communicate_ghost_buffers``,
problem.py:676-707, is a local copy), so this module has no counterpart there.  What it
must preserve is the reference's result: a rank's x-ghost row on a slab-internal side
holds the neighbour's adjacent interior row *with its y ghosts filled* -- exactly what the
reference's x-then-y fill order leaves there -- so ``q`` is bit-identical to the
single-domain run for any number of ranks.  The loop scalars are folded from one
5-double record per rank, gathered in rank order, so every rank advances ``dt`` identically
(maxima are order-independent -> ``dt`` is bit-identical to the single-domain run; the
kinetic-energy sum differs in the last bits, like any re-ordered sum).

Per ``Problem.update()``: 2 passes x (1 neighbour exchange of 3*(Ny+2) doubles per side)
+ 1 all-gather of 5 doubles.  No data-path collective beyond that.
"""
import ctypes as C
import os

import numpy as np
import torch
import torch.distributed as dist

from . import _capi


def slab_rows(Nx, world):
    """[(first interior row (0-based, global), rows)] per rank: as even as possible."""
    base, rem = divmod(Nx, world)
    out, start = [], 0
    for r in range(world):
        n = base + (1 if r < rem else 0)
        out.append((start, n))
        start += n
    return out


def x_periodic(grid):
    """Periodicity of the x sides as the reference gates it (problem.py:681, 688)."""
    e, w = all(grid['bc_xE_P']), all(grid['bc_xW_P'])
    if e != w:
        raise _capi.GapflowError(_capi.GPF_E_INVALID, "slab decomposition needs both x sides periodic or neither")
    return e


class HaloExchanger:
    """Neighbour exchange of ``[2, ...]`` buffers: ``send[0]`` (first interior row) goes to the
    lower neighbour, ``send[1]`` (last interior row) to the upper one; ``recv[0]`` comes from
    the lower neighbour, ``recv[1]`` from the upper one.  Works on CPU (gloo) and CUDA (nccl)
    tensors.  With 2 ranks on a ring both neighbours are the same peer: sends are posted
    upper-first and receives lower-first so that the peer's messages pair up in order."""

    def __init__(self, rank, world, periodic, group=None):
        self.rank, self.world, self.group = rank, world, group
        self.lo = rank - 1 if rank > 0 else (world - 1 if periodic and world > 1 else None)
        self.hi = rank + 1 if rank < world - 1 else (0 if periodic and world > 1 else None)

    def exchange(self, send, recv):
        ops = []
        if self.hi is not None:
            ops.append(dist.P2POp(dist.isend, send[1], self.hi, self.group))
        if self.lo is not None:
            ops.append(dist.P2POp(dist.isend, send[0], self.lo, self.group))
        if self.lo is not None:
            ops.append(dist.P2POp(dist.irecv, recv[0], self.lo, self.group))
        if self.hi is not None:
            ops.append(dist.P2POp(dist.irecv, recv[1], self.hi, self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()          # CUDA: stream dependency only, the host does not block

    def gather(self, rec, out):
        """all-gather of one small record per rank into ``out[world, ...]`` (rank order)."""
        if self.world == 1:
            out.copy_(rec.reshape(out.shape))
        else:
            dist.all_gather_into_tensor(out.view(-1), rec.view(-1), group=self.group)


def local_grid(grid, rank, world):
    """This rank's grid dict: Nx, Lx of the slab; everything else global."""
    start, n = slab_rows(grid['Nx'], world)[rank]
    g = dict(grid)
    g['Nx'] = n
    g['Lx'] = grid['dx'] * n
    return g, start


class SlabSolver:
    """One rank's slab of a global FV problem.  ``topo``/``q``/``extra`` are this rank's padded
    rows ``start .. start+n+1`` of the global padded arrays (shape ``(ncomp, n+2, Ny+2)``)."""

    def __init__(self, grid, geo, prop, numerics, topo, q=None, extra=None, rank=None, world=None, group=None,
                 device=None, exchanger=None, arith='exact', overlap=True, peer=None):
        from .fv import FVSolver
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world is None else world
        self.grid_global = grid
        per = x_periodic(grid)
        self.ex = exchanger or HaloExchanger(self.rank, self.world, per, group)
        g, self.row0 = local_grid(grid, self.rank, self.world)
        if g['Nx'] < 1:
            raise _capi.GapflowError(_capi.GPF_E_INVALID, "more ranks than grid rows")
        halo = (self.ex.lo is not None, self.ex.hi is not None)
        edge = (halo[0] and self.rank == 0, halo[1] and self.rank == self.world - 1)    # periodic ring ends
        self.fv = FVSolver(g, geo, prop, numerics, topo=topo, extra=extra, device=device, halo=halo,
                           halo_edge=edge, arith=arith)
        self.lib, self.h = self.fv.lib, self.fv.h
        dev = self.fv.device
        Ny2 = g['Ny'] + 2
        self.send = torch.zeros((2, 3, Ny2), dtype=torch.float64, device=dev)
        self.recv = torch.zeros((2, 3, Ny2), dtype=torch.float64, device=dev)
        self.rec = torch.zeros(5, dtype=torch.float64, device=dev)
        self.recs = torch.zeros((self.world, 5), dtype=torch.float64, device=dev)
        self._timing, self._ev = False, []
        self._run_mode = 0
        # edge tile rows first, their exchange on a side stream under the interior launch
        self.overlap = overlap and self.world > 1
        self._side = torch.cuda.Stream(device=dev, priority=-1) if self.overlap else None
        # peer=True: halo rows and records go through peer-mapped memory (CUDA IPC over NVLink) written
        # and awaited by the library's own kernels; None: whenever ranks are separate processes on CUDA
        # devices of one node (GPF_SLAB_PEER=0 keeps the grouped NCCL send/recv + all-gather)
        self._comm = None
        if peer is None:
            peer = (exchanger is None and self.world > 1 and dist.is_initialized() and dist.get_backend(group) == 'nccl'
                    and os.environ.get('GPF_SLAB_PEER', '1') != '0')
        if peer:
            self._connect_peers(g['Ny'], group)
        if exchanger is None:
            # Problem.__init__ computes kinetic_energy_old on the _initialize state (problem.py:670)
            self.pre_run(reset_ekin_old=True)
            if q is not None:
                self.set_q(q)
                self.pre_run(reset_ekin_old=False)

    # -- building blocks ---------------------------------------------------------------
    def _s(self):
        return self.fv._stream()

    def _connect_peers(self, Ny, group):
        """Every rank creates its buffer, the IPC handles are all-gathered, every rank maps the others'."""
        n = _capi.IPC_HANDLE_BYTES
        mine = (C.c_ubyte * n)()
        comm = C.c_void_p()
        _capi.check(self.lib.gpf_comm_create(self.world, self.rank, int(Ny), mine, C.byref(comm)))
        # the handles travel by whatever backend the group has (device tensors for nccl, host tensors for gloo:
        # CUDA IPC itself only needs the ranks to be processes of one node, also two processes on one GPU)
        dev = self.fv.device if dist.get_backend(group) == 'nccl' else torch.device('cpu')
        local = torch.tensor(list(mine), dtype=torch.uint8, device=dev)
        parts = [torch.zeros(n, dtype=torch.uint8, device=dev) for _ in range(self.world)]
        dist.all_gather(parts, local, group=group)
        buf = (C.c_ubyte * (self.world * n)).from_buffer_copy(torch.cat(parts).cpu().numpy().tobytes())
        _capi.check(self.lib.gpf_comm_connect(comm, buf))
        self.fused = False
        if os.environ.get('GPF_SLAB_FUSED', '1') != '0' and not self.fv.cfg.visc_field:
            # fused exchange: every rank exports its q buffers (CUDA IPC of the workspace allocation), the neighbours map
            # them, and the pass kernels store their edge rows straight into the neighbours' ghost rows
            nb = _capi.PEER_BLOB_BYTES
            blob = (C.c_ubyte * nb)()
            rc = self.lib.gpf_fv_peer_export(self.h, blob)
            ok = torch.tensor([1 if rc == 0 else 0], dtype=torch.int32, device=dev)
            oks = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(self.world)]
            dist.all_gather(oks, ok, group=group)
            if all(int(o.item()) for o in oks):          # all ranks or none: the step graphs must match
                mine_b = torch.tensor(list(blob), dtype=torch.uint8, device=dev)
                parts = [torch.zeros(nb, dtype=torch.uint8, device=dev) for _ in range(self.world)]
                dist.all_gather(parts, mine_b, group=group)
                blobs = (C.c_ubyte * (self.world * nb)).from_buffer_copy(torch.cat(parts).cpu().numpy().tobytes())
                lo, hi = self._neighbours()
                rc = self.lib.gpf_comm_attach(comm, self.h, blobs, lo, hi)
                # an attach that fails on one rank only (a mapping refused) must not leave the ranks with different
                # step graphs, nor one rank raising while the others wait in a collective: agree, then all or none
                ok = torch.tensor([1 if rc == 0 else 0], dtype=torch.int32, device=dev)
                dist.all_gather(oks, ok, group=group)
                if all(int(o.item()) for o in oks):
                    self.fused = True
                else:
                    why = self.lib.gpf_last_error().decode() if rc != 0 else "another rank could not map its neighbour"
                    self.lib.gpf_comm_detach(comm)
                    if self.rank == 0 or rc != 0:
                        import warnings
                        warnings.warn(f"gapflow_b200 rank {self.rank}: fused slab exchange not available ({why}); "
                                      "the slab step keeps its exchange kernels")
            elif rc != 0 and self.rank == 0:
                import warnings
                warnings.warn("gapflow_b200: the solver workspace cannot be exported over CUDA IPC ("
                              + self.lib.gpf_last_error().decode() + "); the slab step keeps its exchange kernels")
        dist.barrier(group=group)            # nobody writes into a buffer that is still being zeroed
        self._comm = comm

    def close(self):
        if self._comm is not None:
            torch.cuda.synchronize(self.fv.device)
            if dist.is_initialized():
                dist.barrier(group=self.ex.group)
            self.lib.gpf_comm_destroy(self._comm)
            self._comm = None
            if dist.is_initialized():
                # nobody frees a workspace (when the solver is dropped after close()) that a neighbour still has mapped
                dist.barrier(group=self.ex.group)

    def comm_error(self):
        """True if a device-side wait of THIS rank on a peer timed out (synchronises).  The flag also travels to
        every rank with the next record fold, where it stops the run (gpf_fv_scalars.stop, GPF_E_STATE)."""
        if self._comm is None:
            return False
        e = C.c_int()
        _capi.check(self.lib.gpf_comm_error(self._comm, C.byref(e)))
        return bool(e.value)

    def set_comm_timeout(self, seconds):
        """Bound of the device-side waits on a peer (default 30 s): longer than the largest host-side skew
        between ranks (output, GP training on one rank)."""
        if self._comm is not None:
            _capi.check(self.lib.gpf_comm_set_timeout(self._comm, float(seconds)))

    def _check_comm(self):
        """Called at every host synchronisation point: stale halo rows must never reach a result silently."""
        if self.comm_error():
            raise _capi.GapflowError(_capi.GPF_E_STATE, f"rank {self.rank}: a wait on a neighbour rank timed out "
                                     "(a peer died or lagged beyond the communicator's timeout); fields are invalid")

    def set_q(self, q):
        """this rank's padded slab (3, n+2, Ny+2) of the global padded q"""
        self.fv.set_q(q)

    # the five device-side phases (each enqueues on the current stream)
    def _pass(self, corrector, part=_capi.PART_ALL):
        if self._timing:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        _capi.check(self.lib.gpf_fv_pass_part(self.h, corrector, part, self._s()))
        if self._timing:
            e1.record()
            self._ev.append((corrector, part, e0, e1))

    def _pack(self, rel):
        _capi.check(self.lib.gpf_fv_halo_pack(self.h, rel, C.c_void_p(self.send.data_ptr()), self._s()))

    def _unpack(self, rel):
        _capi.check(self.lib.gpf_fv_halo_unpack(self.h, rel, C.c_void_p(self.recv.data_ptr()), self._s()))

    def _record(self, kind):
        _capi.check(self.lib.gpf_fv_reduce_local(self.h, kind, C.c_void_p(self.rec.data_ptr()), self._s()))

    def _finish(self, mode):
        _capi.check(self.lib.gpf_fv_finish(self.h, mode, C.c_void_p(self.recs.data_ptr()), self.world, self._s()))

    def _exchange(self, rel):
        """halo rows of buffer `rel` (0 current, 1 predictor output, 2 corrector output)"""
        if self.world == 1:
            return
        if self._comm is not None:
            lo, hi = self._neighbours()
            _capi.check(self.lib.gpf_fv_halo_send(self.h, self._comm, rel, lo, hi, self._s()))
            _capi.check(self.lib.gpf_fv_halo_recv(self.h, self._comm, rel, self._s()))
            return
        self._pack(rel)
        self.ex.exchange(self.send, self.recv)
        self._unpack(rel)

    def _neighbours(self):
        return (-1 if self.ex.lo is None else self.ex.lo), (-1 if self.ex.hi is None else self.ex.hi)

    def _fold(self, kind, mode):
        if self._comm is not None:
            _capi.check(self.lib.gpf_fv_record_send(self.h, self._comm, kind, self._s()))
            _capi.check(self.lib.gpf_fv_finish_recv(self.h, self._comm, mode, C.c_void_p(self.recs.data_ptr()), self._s()))
            return
        self._record(kind)
        self.ex.gather(self.rec, self.recs)
        self._finish(mode)

    # -- Problem-level operations ----------------------------------------------------------
    def pre_run(self, reset_ekin_old=False):
        """Problem._pre_run on the global field (problem.py:428-443)."""
        self._fold(1, _capi.FIN_PRERUN if reset_ekin_old else _capi.FIN_PRERUN_KEEP_EKIN)

    def fill_ghosts(self):
        """_communicate_ghost_buffers (problem.py:676-707) across slabs: physical sides locally
        (x first, then y), then the slab-internal rows."""
        self.fv.fill_ghosts()
        self._exchange(0)

    def step(self, n=1, until_converged=False):
        """n x Problem.update() (problem.py:509-569); enqueues only.  ``until_converged``: steps become no-ops on
        the device once the run-loop condition (problem.py:401) is false -- on every rank at the same step."""
        uc = 1 if until_converged else 0
        if uc != self._run_mode:
            _capi.check(self.lib.gpf_fv_set_run_mode(self.h, uc))
            self._run_mode = uc
        if self._comm is not None and not self._timing:
            # peer memory: the whole step is the library's own kernels, replayed as one CUDA graph
            lo, hi = self._neighbours()
            _capi.check(self.lib.gpf_fv_slab_step(self.h, self._comm, lo, hi, int(n), C.c_void_p(self.recs.data_ptr()),
                                                  self._s()))
            return
        main = torch.cuda.current_stream(self.fv.device)
        for _ in range(n):
            for c, rel in ((0, 1), (1, 2)):
                if not self.overlap or self._timing:
                    self._pass(c)
                    self._exchange(rel)
                    continue
                # side stream (high priority): the first and the last tile row, then their exchange;
                # main stream: the interior tile rows meanwhile.  The two launches touch disjoint
                # tiles, records and ghost rows, and a pass never reads the buffer it writes.
                self._side.wait_stream(main)
                with torch.cuda.stream(self._side):
                    self._pass(c, _capi.PART_EDGE)
                    self._exchange(rel)
                self._pass(c, _capi.PART_INTERIOR)
                main.wait_stream(self._side)
            self._fold(0, _capi.FIN_STEP)

    def scalars(self):
        self._fold(1, _capi.FIN_REFRESH)
        sc = self.fv.scalars()              # GPF_E_STATE if any rank's wait timed out (record fold carries it)
        self._check_comm()
        return sc

    def update_host(self, q_host, n=1):
        """Host-buffer step of this rank's slab: upload q_host (padded slab, pinned), n x update(),
        download into q_host.  Loop scalars (dt, kinetic_energy_old) are kept, as in
        gpf_fv_update_host."""
        p, dev = C.c_void_p(q_host.data_ptr()), 0
        _capi.check(self.lib.gpf_fv_upload(self.h, _capi.F_Q, p, 3, dev, self._s()))
        self.step(n)
        _capi.check(self.lib.gpf_fv_download(self.h, _capi.F_Q, p, 3, dev, self._s()))
        torch.cuda.current_stream(self.fv.device).synchronize()
        self._check_comm()

    @property
    def mass(self):
        self._fold(1, _capi.FIN_REFRESH)
        torch.cuda.current_stream(self.fv.device).synchronize()
        self._check_comm()
        return float(self.recs[:, 3].sum().item())

    def get_q(self):
        """this rank's padded slab (3, n+2, Ny+2)"""
        q = self.fv.get_q()
        self._check_comm()
        return q

    def gather_q(self):
        """Global padded q on every rank (tests / output staging; host arrays)."""
        loc = torch.from_numpy(self.get_q())
        rows = slab_rows(self.grid_global['Nx'], self.world)
        Ny2 = self.grid_global['Ny'] + 2
        out = np.empty((3, self.grid_global['Nx'] + 2, Ny2))
        parts = [None] * self.world
        dist.all_gather_object(parts, loc.numpy(), group=self.ex.group)
        for r, (start, n) in enumerate(rows):
            lo = 0 if r == 0 else 1                      # physical ghost row only from the edge ranks
            hi = n + 2 if r == self.world - 1 else n + 1
            out[:, start + lo:start + hi] = parts[r][:, lo:hi]
        return out

    def set_timing(self, on):
        """CUDA events (current stream) around every pass kernel launched from now on."""
        self._timing, self._ev = bool(on), []

    def timing(self):
        """(mean predictor ms, mean corrector ms, launches timed)"""
        torch.cuda.current_stream(self.fv.device).synchronize()
        t, n = {0: 0.0, 1: 0.0}, {0: 0, 1: 0}
        for c, part, e0, e1 in self._ev:
            t[c] += e0.elapsed_time(e1)
            n[c] += part != _capi.PART_INTERIOR            # EDGE + INTERIOR are one pass
        mean = lambda c: t[c] / n[c] if n[c] else 0.0
        return mean(0), mean(1), len(self._ev)

    # -- construction from global host arrays (tests, small problems) -----------------------
    @classmethod
    def from_global(cls, d, q_global, topo_global, extra_global=None, rank=None, world=None, group=None, **kw):
        rank = dist.get_rank(group) if rank is None else rank
        world = dist.get_world_size(group) if world is None else world
        start, n = slab_rows(d['grid']['Nx'], world)[rank]
        sl = slice(start, start + n + 2)
        ex = None if extra_global is None else np.ascontiguousarray(np.asarray(extra_global)[..., sl, :])
        q = None if q_global is None else np.ascontiguousarray(q_global[:, sl])
        return cls(d['grid'], d['geometry'], d['properties'], d['numerics'],
                   topo=np.ascontiguousarray(topo_global[:, sl]), q=q,
                   extra=ex, rank=rank, world=world, group=group, **kw)


class InProcessSlabs:
    """All ranks' slabs on ONE GPU in one process, driven phase by phase in lock step with plain
    tensor copies in place of the NCCL calls.  Exercises every device-side piece of the slab path
    (HALO sides, pack/unpack, per-rank records, gathered finish) where only one GPU is at hand."""

    class _NoComm:
        lo = hi = None
        group = None

    def __init__(self, d, q_global, topo_global, world, extra_global=None, split=True, fused=False, arith='exact'):
        per = x_periodic(d['grid'])
        self.world = world
        self.split = split             # passes as EDGE + INTERIOR launches, like SlabSolver's overlapped step
        self.fused = fused             # the fused exchange of gpf_fv_slab_pass over communicators of this process
        self.s = []
        rows = slab_rows(d['grid']['Nx'], world)
        for r in range(world):
            ex = HaloExchanger(r, world, per)          # neighbour table only; never communicates
            self.s.append(SlabSolver.from_global(d, None, topo_global, extra_global, rank=r, world=world,
                                                 exchanger=ex, arith=arith))
        self.d = d
        if fused:
            self._connect_local(d['grid']['Ny'])
        self._fold(1, _capi.FIN_PRERUN)                # kinetic_energy_old of the _initialize state
        for r, (start, n) in enumerate(rows):
            self.s[r].set_q(np.ascontiguousarray(q_global[:, start:start + n + 2]))
        self._fold(1, _capi.FIN_PRERUN_KEEP_EKIN)

    def _connect_local(self, Ny):
        """one communicator per rank, wired to each other and to the neighbours' q buffers inside this process"""
        lib = self.s[0].lib
        self.comms = []
        for r in range(self.world):
            h = (C.c_ubyte * _capi.IPC_HANDLE_BYTES)()
            c = C.c_void_p()
            _capi.check(lib.gpf_comm_create(self.world, r, int(Ny), h, C.byref(c)))
            self.comms.append(c)
        table = (C.c_void_p * self.world)(*[c.value for c in self.comms])
        for r, s in enumerate(self.s):
            _capi.check(lib.gpf_comm_connect_local(self.comms[r], table))
            lo, hi = s._neighbours()
            _capi.check(lib.gpf_comm_attach_local(self.comms[r], s.h, self.s[lo].h if lo >= 0 else None, lo,
                                                  self.s[hi].h if hi >= 0 else None, hi))

    def close(self):
        if self.fused:
            torch.cuda.synchronize()
            for c in self.comms:
                self.s[0].lib.gpf_comm_destroy(c)
            self.comms, self.fused = [], False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _exchange(self, rel):
        for s in self.s:
            s._pack(rel)
        for s in self.s:
            if s.ex.lo is not None:
                s.recv[0].copy_(self.s[s.ex.lo].send[1])
            if s.ex.hi is not None:
                s.recv[1].copy_(self.s[s.ex.hi].send[0])
        for s in self.s:
            s._unpack(rel)

    def _fold(self, kind, mode):
        for s in self.s:
            s._record(kind)
        recs = torch.stack([s.rec for s in self.s])
        for s in self.s:
            s.recs.copy_(recs)
            s._finish(mode)

    def step(self, n=1):
        if self.fused:
            # lock step, pass by pass: every flag a kernel waits for was raised by a kernel launched before it
            lib = self.s[0].lib
            for _ in range(n):
                for c in (0, 1):
                    for r, s in enumerate(self.s):
                        _capi.check(lib.gpf_fv_slab_pass(s.h, self.comms[r], c, s._s()))
                for r, s in enumerate(self.s):
                    _capi.check(lib.gpf_fv_finish_recv(s.h, self.comms[r], _capi.FIN_STEP, C.c_void_p(s.recs.data_ptr()), s._s()))
            return
        for _ in range(n):
            for c, rel in ((0, 1), (1, 2)):
                for s in self.s:
                    if self.split:
                        s._pass(c, _capi.PART_EDGE)
                        s._pass(c, _capi.PART_INTERIOR)
                    else:
                        s._pass(c)
                self._exchange(rel)
            self._fold(0, _capi.FIN_STEP)

    def fill_ghosts(self):
        for s in self.s:
            s.fv.fill_ghosts()
        self._exchange(0)

    def scalars(self):
        self._fold(1, _capi.FIN_REFRESH)
        return [s.fv.scalars() for s in self.s]

    def gather_q(self):
        g = self.d['grid']
        out = np.empty((3, g['Nx'] + 2, g['Ny'] + 2))
        for r, (start, n) in enumerate(slab_rows(g['Nx'], self.world)):
            loc = self.s[r].get_q()
            lo = 0 if r == 0 else 1
            hi = n + 2 if r == self.world - 1 else n + 1
            out[:, start + lo:start + hi] = loc[:, lo:hi]
        return out


# -------------------------------------------------------------------------------------------
# GP inference across ranks: training set (L, alpha) replicated, query cells partitioned
# -------------------------------------------------------------------------------------------
def fold_argmax(vmax, imax, group=None):
    """Global (maximum variance, index of its FIRST occurrence) from each rank's local pair
    (gp.py:406, 424: jnp.max / jnp.argmax over all cells; NaN propagates like jnp.max).
    ``imax`` must already be a GLOBAL cell index.  Works on CPU (gloo) and CUDA (nccl) tensors."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rec = torch.stack([vmax.reshape(()).to(torch.float64), imax.reshape(()).to(torch.float64)])
    if world == 1:
        return float(rec[0]), int(rec[1])
    out = torch.zeros((world, 2), dtype=torch.float64, device=rec.device)
    dist.all_gather_into_tensor(out.view(-1), rec.view(-1), group=group)
    v, i = out[:, 0].cpu().numpy(), out[:, 1].cpu().numpy().astype(np.int64)
    if np.isnan(v).any():
        k = int(np.min(i[np.isnan(v)]))
        return float('nan'), k
    best = v.max()
    return float(best), int(i[v == best].min())


class PartitionedGP:
    """One GP model replicated on every rank; each rank predicts on its slab's cells only.

    ``offset`` is the global flat index (row-major over the global padded grid) of this rank's
    first query row, so the arg-max is the reference's global ``imax`` (gp.py:424).  Slab-internal
    halo rows must not be queried twice: pass the rank's OWNED rows (``owned_rows``)."""

    def __init__(self, model, group=None):
        self.model, self.group = model, group

    def factor(self, *a, **k):
        self.model.factor(*a, **k)           # M^3/3 flops, redundantly on every rank: no broadcast of L needed

    def mean(self, X_local):
        return self.model.repredict_mean(X_local)

    def mean_var(self, X_local, offset, noise=0.0):
        m, v = self.model.repredict_mean_var(X_local, noise)
        vmax = torch.tensor(self.model._vmax_norm, dtype=torch.float64, device=X_local.device)
        imax = torch.tensor(float(self.model.argmax + offset), dtype=torch.float64, device=X_local.device)
        self.maximum_variance, self.argmax = fold_argmax(vmax, imax, self.group)
        return m, v

    def v_sound_gradient_max(self, X_local):
        _, g = self.model.repredict_mean(X_local, with_gradient_max=True)
        t = torch.tensor([g], dtype=torch.float64, device=X_local.device)
        if dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return float(t.item())


def owned_rows(Nx, rank, world):
    """Padded-array rows [a, b) of the GLOBAL grid that rank `rank` owns for cell-wise work
    (every global row exactly once; the edge ranks own the global ghost rows)."""
    start, n = slab_rows(Nx, world)[rank]
    a = start + (0 if rank == 0 else 1)
    b = start + n + (2 if rank == world - 1 else 1)
    return a, b


# -------------------------------------------------------------------------------------------
# Problem-level view of a slab-decomposed run (gapflow_b200.problem.Problem under torchrun)
# -------------------------------------------------------------------------------------------
class Parallel:
    """Where this process stands in a slab-decomposed ``Problem``: rank, world, the host-level collective
    (``allgather`` of small Python objects, rank order) and -- for ranks that are threads of one process --
    the tensor exchanger that stands in for the device-side exchange.  ``Parallel.auto()`` is what ``Problem``
    uses when none is given: the initialised ``torch.distributed`` group, else ``RANK`` / ``WORLD_SIZE`` /
    ``LOCAL_RANK`` of a torchrun launch (NCCL, one GPU per rank), else a single domain."""

    def __init__(self, rank=0, world=1, coll=None, exchanger=None, group=None):
        from .surrogate import LocalColl
        self.rank, self.world, self.group = rank, world, group
        self.coll = coll or LocalColl()
        self._exchanger = exchanger

    def exchanger(self, periodic):
        return None if self._exchanger is None else self._exchanger(self.rank, periodic)

    @classmethod
    def auto(cls):
        if not dist.is_available():
            return cls()
        if not dist.is_initialized():
            if int(os.environ.get('WORLD_SIZE', '1')) <= 1 or 'RANK' not in os.environ:
                return cls()
            local = int(os.environ.get('LOCAL_RANK', os.environ['RANK']))
            torch.cuda.set_device(local)
            dist.init_process_group('nccl', device_id=torch.device('cuda', local))
        if dist.get_world_size() == 1:
            return cls()
        from .surrogate import DistColl
        return cls(dist.get_rank(), dist.get_world_size(), DistColl())

    @classmethod
    def threads(cls, tc, rank):
        """rank ``rank`` of a ``ThreadColl`` (all slabs in one process, one thread per rank, one GPU)."""
        return cls(rank, tc.world, tc.view(rank), exchanger=tc.exchanger)


class SlabField:
    """The ``FVSolver`` interface ``Problem`` drives, on one rank's slab of the global problem: ``q``, the closure
    fields and the scalars are those of the GLOBAL padded arrays (gathered over the ranks on access: output and
    host edits only, never inside the time loop), so ``Problem`` keeps the reference's semantics
    (problem.py:314-362, 616-637) for any number of ranks."""

    def __init__(self, grid, geo, prop, numerics, topo_global, extra_global, par, arith='exact'):
        self.par, self.grid = par, grid
        self.start, self.n = slab_rows(grid['Nx'], par.world)[par.rank]
        sl = slice(self.start, self.start + self.n + 2)
        ex = None if extra_global is None else np.ascontiguousarray(np.asarray(extra_global)[..., sl, :])
        xch = par.exchanger(x_periodic(grid))
        self.slab = SlabSolver(grid, geo, prop, numerics, topo=np.ascontiguousarray(topo_global[:, sl]), extra=ex,
                               rank=par.rank, world=par.world, group=par.group, exchanger=xch, arith=arith)
        if xch is not None:                       # SlabSolver leaves the first fold to a caller that brings its own exchanger
            self.slab.pre_run(reset_ekin_old=True)
        self.fv = self.slab.fv
        self.device, self.lib, self.h = self.fv.device, self.fv.lib, self.fv.h
        self.shape = (grid['Nx'] + 2, grid['Ny'] + 2)           # of the GLOBAL padded arrays
        self.local_shape = self.fv.shape

    # rows of the global padded array this rank contributes when a field is assembled (every row exactly once)
    def _own(self):
        a, b = owned_rows(self.grid['Nx'], self.par.rank, self.par.world)
        return a - self.start, b - self.start

    def gather_rows(self, local):
        """global (..., Nx+2, Ny+2) host array from every rank's local (..., n+2, Ny+2) array (host or device)."""
        if isinstance(local, torch.Tensor):
            local = local.cpu().numpy()
        lo, hi = self._own()
        parts = self.par.coll.allgather(np.ascontiguousarray(local[..., lo:hi, :]))
        return np.concatenate(parts, axis=-2)

    def get_q(self, out=None):
        q = self.gather_rows(self.slab.get_q())
        if out is None:
            return q
        out[...] = q
        return out

    def set_q(self, q):
        self.slab.set_q(np.ascontiguousarray(np.asarray(q)[:, self.start:self.start + self.n + 2]))

    def closures_local(self, stale=False):
        return self.fv.closures(stale=stale)

    def gather_fields(self, fields):
        return tuple(self.gather_rows(f) for f in fields)

    def closures(self, on_device=False, stale=False):
        return self.gather_fields(self.closures_local(stale))

    def pre_run(self, reset_ekin_old=False):
        self.slab.pre_run(reset_ekin_old)

    def step(self, n=1, until_converged=False):
        self.slab.step(n, until_converged)

    def fill_ghosts(self):
        self.slab.fill_ghosts()

    def scalars(self):
        return self.slab.scalars()

    def set_scalars(self, **kw):
        self.fv.set_scalars(**kw)          # the loop state is replicated: every rank applies the same edit

    mass = property(lambda self: self.slab.mass)

    def close(self):
        self.slab.close()
