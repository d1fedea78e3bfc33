"""Device-resident finite-volume solver: thin host mirror over the C ABI.

``FVSolver`` owns nothing but a torch uint8 tensor (the workspace) and an opaque handle;
all arithmetic happens in libgapflow_b200.so.  It mirrors the state and properties of
``GaPFlow.problem.Problem`` that belong to the FV hot path (problem.py:314-362, 509-610).
"""
import ctypes as C

import numpy as np
import torch

from . import _capi


def _ptr(x):
    """Device/host pointer + on_device flag of a torch tensor or NumPy array."""
    if isinstance(x, torch.Tensor):
        assert x.dtype == torch.float64 and x.is_contiguous()
        return C.c_void_p(x.data_ptr()), 1 if x.is_cuda else 0
    assert isinstance(x, np.ndarray) and x.dtype == np.float64 and x.flags.c_contiguous
    return C.c_void_p(x.ctypes.data), 0


class FVSolver:
    def __init__(self, grid, geo, prop, numerics, topo, extra=None, device=None, halo=(False, False),
                 halo_edge=(False, False), arith='exact'):
        if not torch.cuda.is_available():
            raise _capi.GapflowLibraryError("gapflow_b200 needs a CUDA device; there is no CPU path")
        self.lib = _capi.load()
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        torch.cuda.set_device(self.device)
        self.grid, self.geo, self.prop, self.numerics = grid, geo, prop, numerics
        self.shape = (grid['Nx'] + 2, grid['Ny'] + 2)
        has_extra = extra is not None and bool(np.any(np.asarray(extra) != 0.0))
        self.cfg = _capi.make_cfg(grid, geo, prop, numerics, has_extra=has_extra, halo=halo, halo_edge=halo_edge,
                                  arith=arith)
        nbytes = C.c_size_t()
        _capi.check(self.lib.gpf_fv_workspace_bytes(C.byref(self.cfg), C.byref(nbytes)))
        self.workspace = torch.empty(nbytes.value, dtype=torch.uint8, device=self.device)
        h = C.c_void_p()
        _capi.check(self.lib.gpf_fv_create(C.byref(self.cfg), C.c_void_p(self.workspace.data_ptr()),
                                           nbytes.value, C.byref(h)))
        self.h = h
        if self.cfg.visc_field:
            # viscosity modifiers (stress.py:305-326): laws + the cell-centre coordinates np.gradient gets
            # (stress.py:180, 192: topo.x[:, 0], topo.y[0, :]; topography.py:38-54)
            from .topography import midpoint_grid
            xx, yy = midpoint_grid(grid)
            self._xy = (np.ascontiguousarray(xx[:, 0], dtype=np.float64), np.ascontiguousarray(yy[0, :], dtype=np.float64))
            pz, pp, th, tp = _capi.viscosity_args(prop)
            _capi.check(self.lib.gpf_fv_set_viscosity(self.h, pz, pp, th, tp, C.c_void_p(self._xy[0].ctypes.data),
                                                      C.c_void_p(self._xy[1].ctypes.data)))
        self.upload_topo(topo)
        if has_extra:
            e = np.ascontiguousarray(np.broadcast_to(np.asarray(extra, dtype=np.float64), (1, *self.shape)))
            p, dev = _ptr(e)
            _capi.check(self.lib.gpf_fv_upload(self.h, _capi.F_EXTRA, p, 1, dev, self._stream()))
        self.pre_run(reset_ekin_old=True)

    def __del__(self):
        h = getattr(self, 'h', None)
        if h:
            self.lib.gpf_fv_destroy(h)
            self.h = None

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # -- fields ---------------------------------------------------------------------
    def upload_topo(self, topo):
        t = np.ascontiguousarray(topo, dtype=np.float64) if not isinstance(topo, torch.Tensor) else topo
        assert tuple(t.shape[1:]) == self.shape and t.shape[0] in (3, 4)
        p, dev = _ptr(t)
        _capi.check(self.lib.gpf_fv_upload(self.h, _capi.F_TOPO, p, int(t.shape[0]), dev, self._stream()))
        torch.cuda.current_stream(self.device).synchronize()

    def set_q(self, q):
        """Replace the solution field (dense (3, Nx+2, Ny+2), NumPy or torch)."""
        if not isinstance(q, torch.Tensor):
            q = np.ascontiguousarray(q, dtype=np.float64)
        assert tuple(q.shape) == (3, *self.shape)
        p, dev = _ptr(q)
        _capi.check(self.lib.gpf_fv_upload(self.h, _capi.F_Q, p, 3, dev, self._stream()))
        torch.cuda.current_stream(self.device).synchronize()

    def get_q(self, out=None):
        """Solution field as a dense array: NumPy by default, or into ``out`` (torch/NumPy)."""
        if out is None:
            out = np.empty((3, *self.shape))
        p, dev = _ptr(out)
        _capi.check(self.lib.gpf_fv_download(self.h, _capi.F_Q, p, 3, dev, self._stream()))
        torch.cuda.current_stream(self.device).synchronize()
        return out

    def closures(self, on_device=False, stale=False):
        """(pressure, wall_stress_xz, wall_stress_yz, bulk_viscous_stress) of the current q, or with
        ``stale=True`` of the last corrector pass's input: the state the reference's fields are in
        when ``Problem.write()`` stores them (problem.py:532-540, 629)."""
        if on_device:
            mk = lambda n: torch.empty((n, *self.shape), dtype=torch.float64, device=self.device)
        else:
            mk = lambda n: np.empty((n, *self.shape))
        outs = [mk(1), mk(12), mk(12), mk(3)]
        ptrs = [_ptr(o)[0] for o in outs]
        _capi.check(self.lib.gpf_fv_closures_of(self.h, 1 if stale else 0, *ptrs, 1 if on_device else 0, self._stream()))
        torch.cuda.current_stream(self.device).synchronize()
        return outs[0][0], outs[1], outs[2], outs[3]

    # -- time loop --------------------------------------------------------------------
    def viscosity(self, stale=False):
        """The shear-viscosity field (Nx+2, Ny+2) of the current q (``stale``: of the last corrector pass's
        input) when properties has a piezo / thinning block (stress.py:305-326); NumPy array."""
        out = np.empty(self.shape)
        _capi.check(self.lib.gpf_fv_viscosity(self.h, 1 if stale else 0, C.c_void_p(out.ctypes.data), 0, self._stream()))
        return out

    def pre_run(self, reset_ekin_old=False):
        _capi.check(self.lib.gpf_fv_pre_run(self.h, 1 if reset_ekin_old else 0, self._stream()))

    def step(self, n=1, until_converged=False):
        """Enqueue n x Problem.update() (no synchronisation)."""
        _capi.check(self.lib.gpf_fv_step(self.h, int(n), 1 if until_converged else 0, self._stream()))

    def fill_ghosts(self):
        _capi.check(self.lib.gpf_fv_fill_ghosts(self.h, self._stream()))

    def scalars(self):
        s = _capi.FvScalars()
        _capi.check(self.lib.gpf_fv_get_scalars(self.h, C.byref(s), self._stream()))
        return s

    def set_scalars(self, **kw):
        s = self.scalars()
        for k, v in kw.items():
            if k == 'residual_buffer':
                v = list(v)
                s.residual_count = len(v)
                for i, x in enumerate(v):
                    s.residual_buffer[i] = x
            else:
                setattr(s, k, v)
        _capi.check(self.lib.gpf_fv_set_scalars(self.h, C.byref(s), self._stream()))

    def update_host(self, q_host, n=1):
        """Host-buffer step: upload q_host, n x update(), download into q_host."""
        assert isinstance(q_host, (np.ndarray, torch.Tensor))
        p, dev = _ptr(q_host)
        assert dev == 0
        s = _capi.FvScalars()
        _capi.check(self.lib.gpf_fv_update_host(self.h, p, int(n), C.byref(s), self._stream()))
        return s

    @property
    def mass(self):
        m = C.c_double()
        _capi.check(self.lib.gpf_fv_mass(self.h, C.byref(m), self._stream()))
        return m.value

    def set_timing(self, on):
        _capi.check(self.lib.gpf_fv_set_timing(self.h, 1 if on else 0))

    def timing(self):
        a, b, n = C.c_double(), C.c_double(), C.c_int64()
        _capi.check(self.lib.gpf_fv_get_timing(self.h, C.byref(a), C.byref(b), C.byref(n)))
        return a.value, b.value, n.value


# -- pure functions (reference: GaPFlow/models/pressure.py:35, sound.py:35, viscous.py) ----
def _pure_cfg(prop, geo=None):
    grid = dict(Nx=1, Ny=1, dx=1., dy=1.)
    for s in ['xE', 'xW', 'yS', 'yN']:
        grid[f'bc_{s}_P'] = [True] * 3
        grid[f'bc_{s}_D'] = [False] * 3
        grid[f'bc_{s}_N'] = [False] * 3
    geo = geo or dict(U=0., V=0.)
    num = dict(adaptive=False, CFL=0.5, dt=1., tol=0., max_it=1, MC_order=1)
    p = dict(prop)
    p.setdefault('shear', 0.)
    p.setdefault('bulk', 0.)
    p.setdefault('rho0', 1.)
    return _capi.make_cfg(grid, geo, p, num)


def eos_pressure(density, prop):
    """eos_pressure(density, prop) on the device; density: CUDA float64 tensor."""
    lib = _capi.load()
    cfg = _pure_cfg(prop)
    d = density.contiguous()
    out = torch.empty_like(d)
    _capi.check(lib.gpf_eos_pressure(C.byref(cfg), C.c_void_p(d.data_ptr()), C.c_void_p(out.data_ptr()),
                                     d.numel(), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out


def eos_sound_velocity(density, prop):
    lib = _capi.load()
    cfg = _pure_cfg(prop)
    d = density.contiguous()
    out = torch.empty_like(d)
    _capi.check(lib.gpf_eos_sound_velocity(C.byref(cfg), C.c_void_p(d.data_ptr()), C.c_void_p(out.data_ptr()),
                                           d.numel(), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out


def _stress(name, ncomp, q, h, U, V, eta, zeta, Ls, dqx=None, dqy=None, slip="top"):
    """stress_bottom / stress_top / stress_avg(q, h, U, V, eta, zeta, Ls, dqx=None, dqy=None, slip="top") of
    GaPFlow/models/viscous.py:37, 281, 612 on CUDA float64 tensors.  The driver's call (slip="top", no gradients)
    takes the bit-exact kernels; gradients or slip="both" the derived closed forms (gpf_stress_ex, <= 1e-12)."""
    lib = _capi.load()
    if slip not in _capi.SLIP:
        raise ValueError("slip must be 'top' or 'both'")       # viscous.py has no other branch
    cfg = _pure_cfg(dict(EOS='DH', rho0=1., P0=0., C1=1., C2=2., shear=eta, bulk=zeta), dict(U=U, V=V))
    q = q.contiguous()
    h = h[:3].contiguous()
    n = q[0].numel()
    out = torch.empty((ncomp, *q.shape[1:]), dtype=torch.float64, device=q.device)
    lsp = C.c_void_p(0)
    if Ls is not None:
        Ls = torch.broadcast_to(torch.as_tensor(Ls, dtype=torch.float64, device=q.device), q.shape[1:]).contiguous()
        lsp = C.c_void_p(Ls.data_ptr())
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if dqx is None and dqy is None and slip == "top":
        _capi.check(getattr(lib, name)(C.byref(cfg), C.c_void_p(q.data_ptr()), C.c_void_p(h.data_ptr()), lsp,
                                       C.c_void_p(out.data_ptr()), n, stream))
        return out
    gx = None if dqx is None else dqx.contiguous()
    gy = None if dqy is None else dqy.contiguous()
    which = {"gpf_stress_bottom": 0, "gpf_stress_top": 1, "gpf_stress_avg": 2}[name]
    _capi.check(lib.gpf_stress_ex(C.byref(cfg), which, _capi.SLIP[slip], C.c_void_p(q.data_ptr()), C.c_void_p(h.data_ptr()), lsp,
                                  C.c_void_p(gx.data_ptr()) if gx is not None else None,
                                  C.c_void_p(gy.data_ptr()) if gy is not None else None,
                                  C.c_void_p(out.data_ptr()), n, stream))
    return out


def stress_bottom(q, h, U, V, eta, zeta, Ls=None, dqx=None, dqy=None, slip="top"):
    return _stress("gpf_stress_bottom", 6, q, h, U, V, eta, zeta, Ls, dqx, dqy, slip)


def stress_top(q, h, U, V, eta, zeta, Ls=None, dqx=None, dqy=None, slip="top"):
    return _stress("gpf_stress_top", 6, q, h, U, V, eta, zeta, Ls, dqx, dqy, slip)


def stress_avg(q, h, U, V, eta, zeta, Ls=None, dqx=None, dqy=None, slip="top"):
    return _stress("gpf_stress_avg", 3, q, h, U, V, eta, zeta, Ls, dqx, dqy, slip)
