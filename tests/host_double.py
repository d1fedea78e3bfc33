"""A CPU stand-in for ``gapflow_b200.fv.FVSolver`` built on the oracle (tests only, not a test module).

``gapflow_b200.problem.Problem`` is host logic around a device solver: host edits of ``q`` and ``dt``, the run loop and
its batches, when outputs are written and what they hold, the closure fields as the reference keeps them between steps.
With the device replaced by this double that logic runs in the CPU suite, next to the EXECUTED reference ``Problem``
(tests/test_problem_host_cpu.py).  The double implements the part of the FVSolver interface that ``Problem`` calls, with
FVSolver's semantics (gapflow_b200/fv.py, include/gapflow_b200.h: gpf_fv_pre_run, gpf_fv_step with ``until_converged``,
gpf_fv_get_scalars, gpf_fv_closures_of); it is NOT a fallback of the product: nothing in ``gapflow_b200`` imports it."""
import types

import numpy as np

from oracle.fv import FVOracle


class OracleFVSolver:
    def __init__(self, grid, geo, prop, numerics, topo=None, extra=None, arith='exact', **kw):
        self.o = FVOracle(grid, geo, prop, numerics, extra=extra, topo=topo)
        self.numerics = numerics
        self.shape = self.o.q.shape[1:]
        self.stop = False
        self.calls = []                      # (name, args) of the calls Problem makes, for the tests

    # -- fields -----------------------------------------------------------------------------------------------------
    def get_q(self, out=None):
        self.calls.append(('get_q',))
        if out is None:
            return self.o.q.copy()
        out[...] = self.o.q
        return out

    def set_q(self, q):
        self.calls.append(('set_q',))
        self.o.q[...] = q

    def fill_ghosts(self):
        self.calls.append(('fill_ghosts',))
        self.o.fill_ghosts()

    mass = property(lambda self: self.o.mass)

    def closures(self, stale=False):
        """gpf_fv_closures_of: stale -> the fields of the corrector pass's input (what the reference's muGrid fields hold
        between steps), else evaluated on the current q"""
        self.calls.append(('closures', stale))
        o = self.o
        if not stale:
            o.closures()
        return tuple(np.array(a) for a in (o.pressure, o.wall_xz, o.wall_yz, o.bulk))

    # -- loop state ---------------------------------------------------------------------------------------------------
    def pre_run(self, reset_ekin_old=False):
        self.calls.append(('pre_run', reset_ekin_old))
        o = self.o
        o.step, o.simtime, o.residual = 0, 0., 1.
        o.residual_buffer.clear()
        o.residual_buffer.append(1.)
        o.dt = self.numerics['CFL'] * o.dt_crit if self.numerics['adaptive'] else self.numerics['dt']
        if reset_ekin_old:
            o.kinetic_energy_old = o.kinetic_energy
        self.stop = o._stop = False

    def scalars(self):
        o = self.o
        buf = list(o.residual_buffer)
        with np.errstate(all='ignore'):
            return types.SimpleNamespace(
                dt=o.dt, step=o.step, simtime=o.simtime, residual=o.residual, residual_buffer=buf + [0.] * (5 - len(buf)),
                residual_count=len(buf), ekin=o.kinetic_energy, ekin_old=o.kinetic_energy_old, v_max=o.v_max,
                v_sound=o.v_sound, dt_crit=o.dt_crit, cfl=o.cfl, has_nan=int(np.any(np.isnan(o.q))),
                has_neg=int(np.any(o.q[0] < 0.)), stop=int(o._stop or self.stop), converged=int(o.converged))

    def set_scalars(self, **kw):
        self.calls.append(('set_scalars', tuple(sorted(kw))))
        for k, v in kw.items():
            if k == 'dt':
                self.o.dt = v
            elif k == 'ekin_old':
                self.o.kinetic_energy_old = v
            elif k == 'stop':
                self.stop = self.o._stop = bool(v)
            else:
                raise KeyError(k)

    def step(self, n=1, until_converged=False):
        """gpf_fv_step: with until_converged the loop condition of Problem.run (problem.py:401) is evaluated before
        every step; steps past it are no-ops"""
        self.calls.append(('step', n, until_converged))
        o = self.o
        for _ in range(n):
            if o._stop or (until_converged and (o.converged or o.step >= o.max_it)):
                break
            o.update()


# ---------------------------------------------------------------------------------------------------------------------
# GP-coupled step: doubles for the device pieces gapflow_b200/surrogate.py drives (the pass kernels with GP closures, the
# feature kernel, the GP model).  SurrogateDriver / SurrogateModel / Database -- the host logic -- run unchanged.
# ---------------------------------------------------------------------------------------------------------------------
import ctypes as C  # noqa: E402

import torch  # noqa: E402

from oracle import closures as OC  # noqa: E402
from oracle import gp as OG  # noqa: E402
from oracle.fv import predictor_corrector, source  # noqa: E402


class _SurrogateOracle(FVOracle):
    """FVOracle whose sound speed is the one handed in with the step's fold (stress.py:522-537 evaluated by the driver)"""
    c2 = None

    @property
    def v_sound(self):
        if self.c2 is not None:
            return np.sqrt(self.c2)
        return OC.eos_sound_velocity(self.q[0], self.prop).max()


class OracleSurrogateSolver(OracleFVSolver):
    """OracleFVSolver + what SurrogateDriver asks of the device: gpf_fv_set_surrogate, gpf_fv_pass, gpf_gp_features_of,
    gpf_fv_reduce_local / gpf_fv_finish_surrogate (reached through FakeLib below)."""

    def __init__(self, grid, geo, prop, numerics, topo=None, extra=None, arith='exact', **kw):
        super().__init__(grid, geo, prop, numerics, topo=topo, extra=extra, arith=arith)
        self.o.__class__ = _SurrogateOracle
        self.h, self.device = self, torch.device('cpu')
        self.bound = {}
        self.stage = {}

    def _stream(self):
        return None

    def field(self, rel):
        """the q of rotating buffer `rel`: 0 the step's input, 1 the predictor's output, 2 the new field"""
        return self.stage.get(rel, self.o.q) if rel else (self.stage.get(0, self.o.q) if self._in_step else self.o.q)

    _in_step = False

    def run_pass(self, i):
        o = self.o
        if i == 0:
            self._in_step = True
            self.stage = {0: o.q.copy()}
            self._dt = o.dt
        FVOracle.closures(o)                                   # fixed-form everything, then the GP means on top
        n = self.shape[0] * self.shape[1]
        for key, (ptr, scale) in self.bound.items():
            if ptr is None:
                continue
            nout = 1 if key == 'zz' else 2
            m = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(nout * n,)).reshape(nout, *self.shape) * scale
            if key == 'zz':
                o.pressure[...] = m[0]
            else:
                f, oi = (o.wall_xz, 4) if key == 'xz' else (o.wall_yz, 3)
                f[oi], f[oi + 6] = m[0], m[1]
        d = o.directions()[i]
        fX, fY = predictor_corrector(o.q, o.pressure, o.bulk, d)
        src = source(o.q, o.topo, o.bulk, o.wall_xz[:6] + o.wall_yz[:6], o.wall_xz[6:] + o.wall_yz[6:])
        o.q[...] = o.q - self._dt * (fX / o.grid['dx'] + fY / o.grid['dy'] - src)
        o.fill_ghosts()
        if i == 0:
            self.stage[1] = o.q
        else:
            o.q[...] = (o.q + self.stage[0]) / 2.0
            self._valid = o.q_is_valid
            if self._valid:
                o.fill_ghosts()
            self.stage[2] = o.q

    def finish(self, mode, c2):
        from gapflow_b200 import _capi
        o = self.o
        o.c2 = c2 if c2 > 0 or c2 != c2 else None
        if mode == _capi.FIN_STEP:
            self._in_step = False
            if self._valid:
                ekin = o.kinetic_energy
                o.residual = abs(ekin - o.kinetic_energy_old) / o.kinetic_energy_old / o.cfl
                o.residual_buffer.append(o.residual)
                o.kinetic_energy_old = ekin
                o.step += 1
                o.simtime += o.dt
                if self.numerics['adaptive']:
                    o.dt = self.numerics['CFL'] * o.dt_crit
            else:
                o.q[...] = self.stage[0]
                o._stop = True
        elif mode in (_capi.FIN_PRERUN, _capi.FIN_PRERUN_KEEP_EKIN):
            self.pre_run(reset_ekin_old=mode == _capi.FIN_PRERUN)


class FakeLib:
    """the entry points of libgapflow_b200.so that gapflow_b200/surrogate.py calls, on the doubles above"""

    def gpf_fv_set_surrogate(self, h, p, ps, xz, xs, yz, ys):
        h.bound = {'zz': (p, ps), 'xz': (xz, xs), 'yz': (yz, ys)}
        return 0

    def gpf_fv_pass(self, h, i, stream):
        h.run_pass(i)
        return 0

    def gpf_fv_reduce_local(self, h, kind, rec, stream):
        return 0

    def gpf_fv_finish_surrogate(self, h, mode, rec, world, c2, stream):
        h.finish(mode, c2)
        return 0


class CapiShim:
    """gapflow_b200._capi as surrogate.py sees it, with load() -> FakeLib"""

    def __init__(self):
        from gapflow_b200 import _capi
        self._real, self._lib = _capi, FakeLib()

    def load(self):
        return self._lib

    def check(self, rc):
        assert rc == 0

    def __getattr__(self, k):
        return getattr(self._real, k)


def features(solver, rel, X_scale, active_dims):
    """gpf_gp_features_of: (Np, D) normalised, column-selected features of rotating buffer `rel` (gp.py:223-232)"""
    o = solver.o
    F = np.vstack([solver.field(rel), o.topo[:3], o.extra]).reshape(7, -1).T
    return torch.from_numpy(np.ascontiguousarray((F / np.asarray(X_scale, dtype=float).reshape(7))[:, [int(d) for d in active_dims]]))


class OracleGPModel:
    """gapflow_b200.gp.GPModel on oracle/gp.py"""

    def __init__(self, M_max, D, n_out, device=None):
        self.M_max, self.D, self.n_out, self.M = M_max, D, n_out, 0

    def factor(self, X, Y, log_amp, log_scale, yerr):
        self.X, self.log_amp, self.log_scale = np.array(X, dtype=float), float(log_amp), np.array(log_scale, dtype=float)
        self.M = len(self.X)
        self.L = OG.build(self.X, yerr, self.log_amp, self.log_scale)
        self.alpha = OG.alpha(self.L, Y)

    def train(self, X, Y, yerr, log_amp0, log_scale0):
        la, ls, _ = OG.train(X, Y, yerr, log_amp0, log_scale0)
        self.factor(X, Y, la, ls, yerr)

    def repredict_mean(self, X, with_gradient_max=False):
        X = X.numpy()
        mean = torch.from_numpy(np.ascontiguousarray(OG.repredict_mean(self.X, self.log_amp, self.log_scale, self.alpha, X)))
        if with_gradient_max:
            g = OG.dmean_dx0(self.X, self.log_amp, self.log_scale, self.alpha, X)
            return mean, (float('nan') if np.isnan(g).any() else float(g.max()))
        return mean

    def repredict_mean_var(self, X, noise=0.0):
        mean, var = OG.repredict_mean_var(self.X, self.log_amp, self.log_scale, self.L, self.alpha, noise, X.numpy())
        self.argmax, self._vmax_norm = int(np.argmax(var)), float(np.max(var))
        return torch.from_numpy(np.ascontiguousarray(mean)), torch.from_numpy(np.ascontiguousarray(var))
