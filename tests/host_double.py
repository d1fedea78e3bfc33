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
