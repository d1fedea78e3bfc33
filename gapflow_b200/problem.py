"""Drop-in ``Problem`` for the fixed-form path: same constructor chain, attributes and files as
``GaPFlow.problem.Problem`` (problem.py:56-881), with ``update()`` executed by the CUDA library.

Kept from the reference: ``from_yaml`` / ``from_string`` / ``_from_dict`` (problem.py:251-308),
``run(keep_open)`` (368-410), ``_pre_run`` (412-443), ``update`` (509-569), ``write`` (616-637),
``_post_run`` (453-503); attributes ``q`` (writable in place, as tests do: tests/test_wave_decay.py:101),
``dt``, ``step``, ``simtime``, ``residual``, ``residual_buffer``, ``kinetic_energy_old``, ``_stop``;
properties ``mass``, ``kinetic_energy``, ``v_max``, ``dt_crit``, ``cfl``, ``converged``; sub-objects
``pressure.pressure`` / ``.v_sound``, ``wall_stress_xz|yz.lower/.upper/.full``, ``bulk_stress.stress``,
``topo.full/.h/.x/.y``; outputs ``config.yml``, ``topo.nc``, ``sol.nc``, ``history.csv``.

Device residency: ``q`` lives on the GPU.  Reading ``problem.q`` copies it to a host array and marks it
*exposed*: the next ``update()`` uploads that array again, so in-place edits by user code are honoured;
a ``run()`` never exposes it, so the loop runs without host traffic: ``write_freq`` updates are enqueued
at once and the loop condition (problem.py:401) is evaluated on the device.

GP models (``gp`` + ``database`` arguments, or ``gp`` / ``db`` sections of the YAML with the mock MD): the step
is driven pass by pass by ``gapflow_b200.surrogate.SurrogateDriver`` (GP means on the device, active learning
and the optimiser iteration on the host); outputs ``gp_zz.csv`` / ``gp_xz.csv`` / ``gp_yz.csv`` / ``gp_*.txt`` and
``Xtrain.npy`` / ``Ytrain.npy`` / ``Ytrain_err.npy`` as the reference writes them (problem.py:486-503, db.py:268-276).

Not here (host-side): LAMMPS MD runners, dtool storage, elastic deformation, plotting.
NetCDF: muGrid's ``FileIONetCDF`` is not available; ``sol.nc`` /
``topo.nc`` are written as NetCDF-3 with ``scipy.io.netcdf_file`` using the reference's field names
(the exact dimension names of muGrid's files are unverified).
"""
import csv
import io as _io
import os
import signal
from collections import deque
from datetime import datetime

import numpy as np
import yaml

from . import config
from .fv import FVSolver
from .topography import make_topography, midpoint_grid


class _Topo:
    def __init__(self, grid, geo):
        self.full = make_topography(grid, geo)
        self.x, self.y = midpoint_grid(grid)

    @property
    def h(self):
        return self.full[0]

    def update(self):            # elastic deformation is out of scope: no-op (topography.py:257-271)
        pass


class _Closure:
    """Read-only view of one closure field, fetched from the device on access."""

    def __init__(self, problem, index, name):
        self._p, self._i, self.name = problem, index, name
        self.geo, self.prop = problem.geo, problem.prop
        self.is_gp_model = False

    def _get(self):
        return self._p._closures()[self._i]


class _Pressure(_Closure):
    @property
    def pressure(self):
        return self._get()

    @property
    def v_sound(self):
        return self._p._scalars().v_sound

    def write(self):
        pass


class _WallStress(_Closure):
    @property
    def full(self):
        return self._get()

    @property
    def lower(self):
        return self._get()[:6]

    @property
    def upper(self):
        return self._get()[6:]

    def write(self):
        pass


class _BulkStress(_Closure):
    @property
    def stress(self):
        return self._get()


class Problem:
    def __init__(self, options, grid, numerics, prop, geo, gp=None, database=None, extra=None, extra_field=None,
                 parallel=None, arith=None):
        """``parallel``: a ``gapflow_b200.dist.Parallel`` (default ``Parallel.auto()``: the torch.distributed group or
        torchrun's RANK / WORLD_SIZE; one process per GPU).  With more than one rank the grid is slab-decomposed
        along x; every attribute below keeps its GLOBAL meaning and rank 0 writes the files.
        ``arith``: 'exact' (default; bit-identical to the reference's NumPy expressions) or 'fast' (FMA contraction,
        1e-12 per step); also ``numerics['arith']`` of the YAML."""
        if extra is None:
            extra = extra_field                                        # the reference's keyword (problem.py:85)
        if gp is not None and database is None:
            raise ValueError("a GP configuration needs a Database (gapflow_b200.surrogate.Database)")
        if database is not None and not database.has_mock_md:        # problem.py:110-113
            prop['shear'] = 0.
            prop['bulk'] = 0.
        if prop.get('elastic', {}).get('enabled', False):
            raise NotImplementedError("elastic deformation (ContactMechanics) stays on the host")
        self.options, self.grid, self.numerics, self.prop, self.geo = options, grid, numerics, prop, geo
        self.topo = _Topo(grid, geo)
        shape = (grid['Nx'] + 2, grid['Ny'] + 2)
        self._extra = None if extra is None else np.broadcast_to(np.asarray(extra, dtype=float), (1, *shape)).copy()
        from .dist import Parallel, SlabField
        self._par = parallel if parallel is not None else Parallel.auto()
        self._root = self._par.rank == 0
        arith = arith or numerics.get('arith', 'exact')
        self.has_gp_model = gp is not None
        self._sur = None
        if self._par.world == 1:
            self._solver = FVSolver(grid, geo, prop, numerics, topo=self.topo.full, extra=self._extra, arith=arith)
            if gp is not None:
                from .surrogate import SurrogateDriver
                self._sur = SurrogateDriver(self._solver, gp, database, grid['dim'], write_freq=options['write_freq'])
        else:
            # x-slabs, one per rank (problem.py has no counterpart: the reference is single-process)
            self._solver = SlabField(grid, geo, prop, numerics, self.topo.full, self._extra, self._par, arith=arith)
            if gp is not None:
                from .surrogate import SlabSurrogateDriver
                self._sur = SlabSurrogateDriver(self._solver.slab, gp, database, grid['dim'], self._par.coll,
                                                write_freq=options['write_freq'])
        self.database = database
        self._q_host = np.empty((3, *shape))
        self._q_synced = False       # _q_host equals the device field
        self._q_exposed = False      # user code may have edited _q_host
        self._dt_override = None
        self._ekin_old = None
        self.pressure = _Pressure(self, 0, 'zz')
        self.wall_stress_xz = _WallStress(self, 1, 'xz')
        self.wall_stress_yz = _WallStress(self, 2, 'yz')
        self.bulk_stress = _BulkStress(self, 3, 'bulk')
        self.step = None
        self.simtime = 0.
        self.residual = 1.
        self.residual_buffer = deque([1.], 5)
        self.tol, self.max_it = numerics['tol'], numerics['max_it']
        self._stop = False
        self._clos_cache = None
        self.history = {"step": [], "time": [], "ekin": [], "residual": [], "vsound": []}
        self.outdir, self._sol, self._frames = None, None, 0
        if not options['silent']:
            if self._root:
                self._open_output()
            self.outdir = self._par.coll.allgather(self.outdir)[0]       # every rank knows where rank 0 writes

    # -- constructors (problem.py:251-308) --------------------------------------------------
    @classmethod
    def from_yaml(cls, fname):
        with open(fname, 'r') as f:
            return cls._from_dict(config.read_yaml_input(f))

    @classmethod
    def from_string(cls, ymlstring):
        with _io.StringIO(ymlstring) as f:
            return cls._from_dict(config.read_yaml_input(f))

    @classmethod
    def _from_dict(cls, input_dict):
        d = input_dict
        # problem.py:207-249: a db section builds the Database around the mock MD (md section absent) --
        # LAMMPS runners (md section present) stay with the reference
        gp, db, database = d.get('gp'), d.get('db'), None
        if db is not None:
            if d.get('md') is not None:
                raise NotImplementedError("LAMMPS MD runners are out of scope (host-side, GaPFlow/md); pass a "
                                          "Database built around your own runner to Problem(...)")
            from .surrogate import Database, Mock
            database = Database(Mock(d['properties'], d['geometry'], gp), db)
        return cls(d['options'], d['grid'], d['numerics'], d['properties'], d['geometry'],
                   gp=gp, database=database)

    # -- state ------------------------------------------------------------------------------
    @property
    def q(self):
        self._push_dt()
        if not self._q_synced:
            self._solver.get_q(self._q_host)
            self._q_synced = True
        self._q_exposed = True
        return self._q_host

    # problem.py:319-331 (over the padded array, ghost cells included)
    q_has_nan = property(lambda self: bool(np.any(np.isnan(self.q))))
    q_has_negative_density = property(lambda self: bool(np.any(self.q[0] < 0.)))
    q_is_valid = property(lambda self: not self.q_has_nan and not self.q_has_negative_density)

    def _communicate_ghost_buffers(self):
        """problem.py:676-707 on the current field: scripts call it after writing into ``q``."""
        self._sync_to_device()
        self._solver.fill_ghosts()
        self._q_synced, self._clos_cache, self._scal_stale = False, None, True

    def _sync_to_device(self):
        if self._q_exposed:
            self._solver.set_q(self._q_host)
            self._q_exposed = False
            self._clos_cache = None
            self._scal_stale = True          # dt_crit / cfl / v_sound of an edited q: see _scalars

    @property
    def dt(self):
        return self._dt_override if self._dt_override is not None else self._solver.scalars().dt

    @dt.setter
    def dt(self, value):
        self._dt_override = float(value)

    @property
    def kinetic_energy_old(self):
        return self._ekin_old

    @kinetic_energy_old.setter
    def kinetic_energy_old(self, value):
        """tests and scripts assign it after editing q (tests/test_wave_decay.py:102): the residual of
        the next step is taken against it (problem.py:578)."""
        self._ekin_old = float(value)
        self._solver.set_scalars(ekin_old=self._ekin_old)

    def _push_dt(self):
        if self._dt_override is not None:
            self._solver.set_scalars(dt=self._dt_override)
            self._dt_override = None

    def _scalars(self):
        self._sync_to_device()
        self._push_dt()
        if getattr(self, '_scal_stale', False) and self._sur is not None and self._sur.ready and 'zz' in self._sur.models:
            # a pressure GP defines the sound speed (stress.py:522-537): after a host edit of q the scalars of the
            # current field are folded with the slope of the GP mean, not with the fixed-form law
            self._sur.refresh_scalars()
        self._scal_stale = False
        return self._solver.scalars()

    # GP models of the reference's sub-objects (problem.py:141-146); None for a fixed-form law
    gp_models = property(lambda self: {} if self._sur is None else self._sur.models)

    def _pull(self, sc=None):
        sc = sc or self._solver.scalars()
        self.step, self.simtime, self.residual = int(sc.step), sc.simtime, sc.residual
        self.residual_buffer = deque([sc.residual_buffer[k] for k in range(sc.residual_count)], 5)
        self._ekin_old = sc.ekin_old
        if sc.stop:
            self._stop = True
        return sc

    def _closures(self):
        if not self.step and not self._stop:
            # The reference's stress models fill their fields in update() only (stress.py:289-362, 427-459, 600-622):
            # until the first step they hold the zeros they were created with, and frame 0 of sol.nc is written that way
            # (problem.py:397, 629).  (_finalize evaluates them, problem.py:604-607: an invalid FIRST step shows closures.)
            nx, ny = self.grid['Nx'] + 2, self.grid['Ny'] + 2
            return tuple(np.zeros((n, nx, ny)) if n else np.zeros((nx, ny)) for n in (0, 12, 12, 3))
        if self._clos_cache is None:
            self._sync_to_device()
            if self._par.world == 1:
                self._clos_cache = self._solver.closures(stale=bool(self.step))
                if self._sur is not None and self._sur.ready:
                    self._clos_cache = self._sur.closures(self._clos_cache)
            else:                                   # overlay the GP means on this rank's rows, then assemble the global fields
                loc = self._solver.closures_local(stale=bool(self.step))
                if self._sur is not None and self._sur.ready:
                    loc = self._sur.closures(loc)
                self._clos_cache = self._solver.gather_fields(loc)
        return self._clos_cache

    mass = property(lambda self: (self._sync_to_device(), np.float64(self._solver.mass))[1])
    kinetic_energy = property(lambda self: self._scalars().ekin)
    v_max = property(lambda self: self._scalars().v_max)
    dt_crit = property(lambda self: self._scalars().dt_crit)
    cfl = property(lambda self: self._scalars().cfl)

    @property
    def converged(self):
        return bool(np.all(np.array(self.residual_buffer) < self.tol))

    # -- time loop ----------------------------------------------------------------------------
    def _pre_run(self):
        """problem.py:412-443 (fixed-form branch)."""
        self._sync_to_device()
        if self._sur is not None:
            if self.outdir is not None and self._root:
                self.database.output_path = self.outdir
            self._sur.pre_run()
            if not self.options['silent']:
                for m in self._sur.models.values():
                    m.write()
        else:
            self._solver.pre_run(reset_ekin_old=False)
        self._pull()

    def update(self):
        """One MacCormack step (problem.py:509-569)."""
        if self.step is None:
            self._pre_run()
        self._sync_to_device()
        self._push_dt()
        if self._sur is not None:
            self._sur.update(self.step)
        else:
            self._solver.step(1)
        self._q_synced, self._clos_cache, self._scal_stale = False, None, False
        sc = self._pull()
        if sc.stop and not self.options['silent'] and self._root:
            print("Stopping: NaN or negative density detected (state of the previous step restored).")

    def run(self, keep_open=False):
        if self.step is None:
            self._pre_run()
        self._stop = False
        self._solver.set_scalars(stop=0)
        self.history = {"step": [], "time": [], "ekin": [], "residual": [], "vsound": []}
        silent = self.options['silent']
        if not silent:
            if self._root:
                print(61 * '-')
                print(f"{'Step':6s} {'Timestep':10s} {'Time':10s} {'CFL':10s} {'Residual':10s}")
                print(61 * '-')
            self.write(params=False)
        import threading
        old = {}
        if threading.current_thread() is threading.main_thread():     # signal handlers belong to the main thread
            old = {s: signal.signal(s, self._receive_signal) for s in (signal.SIGINT, signal.SIGTERM)}
        self._tic = datetime.now()
        wf = max(1, int(self.options['write_freq']))
        try:
            while not self.converged and self.step < self.max_it and not self._stop:
                self._sync_to_device()
                self._push_dt()
                if self._sur is not None:
                    # every pass needs fresh GP means (and, with active learning, host decisions)
                    self._sur.update(self.step)
                else:
                    n = wf - self.step % wf                       # up to the next multiple of write_freq
                    self._solver.step(n, until_converged=True)    # loop condition evaluated on the device
                self._q_synced, self._clos_cache = False, None
                self._pull()
                if self._par.world > 1:              # a termination signal reaches one rank: all ranks leave together
                    self._stop = any(self._par.coll.allgather(bool(self._stop)))
                if self.step % wf == 0 and not silent:
                    self.write()
        finally:
            for s, h in old.items():
                signal.signal(s, h)
        if not keep_open:
            self._post_run()

    def _receive_signal(self, signum, frame):
        self._stop = True

    def _post_run(self):
        walltime = datetime.now() - self._tic
        silent = self.options['silent']
        if self.step % max(1, int(self.options['write_freq'])) != 0 and not silent:
            self.write()
        if not silent and self._sol is not None:
            self._sol.close()
            self._sol = None
        speed = self.step / max(walltime.total_seconds(), 1e-12)
        if self._root:
            print(33 * '=')
            print("Total walltime   : ", str(walltime).split('.')[0])
            print(f"({speed:.2f} steps/s)")
            print(33 * '=')
        if not silent and self._root:
            with open(os.path.join(self.outdir, 'history.csv'), 'w', newline='') as f:
                w = csv.writer(f)
                w.writerow(list(self.history))
                w.writerows(zip(*self.history.values()))
            for k, m in self.gp_models.items():               # problem.py:486-503: gp_zz.csv, gp_xz.csv, gp_yz.csv
                with open(os.path.join(self.outdir, f'gp_{k}.csv'), 'w', newline='') as f:
                    w = csv.writer(f)
                    w.writerow(list(m.history))
                    w.writerows(zip(*m.history.values()))
                with open(os.path.join(self.outdir, f'gp_{k}.txt'), 'w') as f:
                    print(f"GaussianProcess(kernel={m.kernel_variance!r} * Linear(scale={m.kernel_lengthscale!r}, "
                          f"Matern32(L2Distance)), M={m.gp.M}, diag={m.Yerr**2!r})", file=f)

    # -- output (problem.py:154-205, 616-637) ---------------------------------------------------
    def _open_output(self):
        from scipy.io import netcdf_file
        name, ts = self.options['output'], self.options['use_tstamp']
        stamp = datetime.now().replace(microsecond=0).strftime("%Y-%m-%d_%H%M%S") + '_' if ts else ''
        self.outdir = os.path.join(os.path.dirname(name), stamp + os.path.basename(name))
        if not os.path.exists(self.outdir):
            os.makedirs(self.outdir)
        elif len(os.listdir(self.outdir)) > 0:
            raise RuntimeError('Output path exists and is not empty.')
        with open(os.path.join(self.outdir, 'config.yml'), 'w') as f:
            yaml.dump(dict(version='gapflow_b200', options=self.options, grid=self.grid, numerics=self.numerics,
                           geo=self.geo, prop=self.prop), f)
        # Layout as muGrid's FileIONetCDF leaves it for the reference's readers (GaPFlow/viz/plotting.py:235-240,
        # 307-312, 331-349 index the arrays by position): multi-component fields (frame, component, sub-point = 1,
        # x, y), scalar fields (frame, x, y).  Variable names are the reference's (problem.py:185-205); the
        # dimension NAMES are this writer's (muGrid's are unverified here: INTEGRATION.md).
        nx, ny = self.grid['Nx'] + 2, self.grid['Ny'] + 2
        with netcdf_file(os.path.join(self.outdir, 'topo.nc'), 'w') as t:
            for d, n in (('frame', None), ('comp4', 4), ('subpt', 1), ('nx', nx), ('ny', ny)):
                t.createDimension(d, n)
            v = t.createVariable('topography', 'f8', ('frame', 'comp4', 'subpt', 'nx', 'ny'))
            v[0, :, 0] = self.topo.full
        self._sol = netcdf_file(os.path.join(self.outdir, 'sol.nc'), 'w')
        for d, n in (('frame', None), ('comp3', 3), ('comp12', 12), ('subpt', 1), ('nx', nx), ('ny', ny)):
            self._sol.createDimension(d, n)
        self._sol.createVariable('solution', 'f8', ('frame', 'comp3', 'subpt', 'nx', 'ny'))
        self._sol.createVariable('pressure', 'f8', ('frame', 'nx', 'ny'))
        self._sol.createVariable('wall_stress_xz', 'f8', ('frame', 'comp12', 'subpt', 'nx', 'ny'))
        self._sol.createVariable('wall_stress_yz', 'f8', ('frame', 'comp12', 'subpt', 'nx', 'ny'))
        # predictive-variance fields of the GP models, in the reference's order (problem.py:195-203)
        for key, name in (('xz', 'wall_stress_xz_var'), ('yz', 'wall_stress_yz_var'), ('zz', 'pressure_var')):
            if key in self.gp_models:
                self._sol.createVariable(name, 'f8', ('frame', 'nx', 'ny'))

    def _variance_fields(self):
        """{name: (Nx+2, Ny+2) host array} of the GP models' variance fields as the reference's muGrid fields hold
        them at write time: the variance of the last inference that computed one (stress.py:353-354, 617-618),
        zeros before the first."""
        out = {}
        for key, name in (('xz', 'wall_stress_xz_var'), ('yz', 'wall_stress_yz_var'), ('zz', 'pressure_var')):
            m = self.gp_models.get(key)
            if m is not None:
                v = m.variance if (self.step or self._stop) else None      # the FIELD is filled in update() only, like the closures
                out[name] = np.zeros((self.grid['Nx'] + 2, self.grid['Ny'] + 2)) if v is None else self._gather_field(v)
        return out

    def _gather_field(self, dev_field):
        """a device field over this rank's padded rows -> the global host array"""
        return dev_field.cpu().numpy() if self._par.world == 1 else self._solver.gather_rows(dev_field)

    def write(self, scalars=True, fields=True, params=True):
        """problem.py:616-637."""
        if self.step is None:
            self._pre_run()
        if scalars:
            sc = self._scalars()
            if self._root:
                print(f"{self.step:<6d} {sc.dt:.4e} {self.simtime:.4e} {sc.cfl:.4e} {self.residual:.4e}")
            for k, v in zip(self.history, (self.step, self.simtime, sc.ekin, self.residual, sc.v_sound)):
                self.history[k].append(v)
        if fields and (self._sol is not None or (self._par.world > 1 and not self.options['silent'])):
            p, wxz, wyz, _ = self._closures()             # slabs: gathers, every rank takes part
            if not self._q_synced:
                self._solver.get_q(self._q_host)
                self._q_synced = True
            var = self._variance_fields()
        if fields and self._sol is not None:
            f = self._frames
            self._sol.variables['solution'][f, :, 0] = self._q_host
            self._sol.variables['pressure'][f] = p
            self._sol.variables['wall_stress_xz'][f, :, 0] = wxz
            self._sol.variables['wall_stress_yz'][f, :, 0] = wyz
            for name, v in var.items():
                self._sol.variables[name][f] = v
            self._frames += 1
            self._sol.flush()
        if params:
            # Pressure.write / WallStress.write (gp.py:242-253): one row of hyper-parameter history per GP model
            for key in ('zz', 'xz', 'yz'):
                m = self.gp_models.get(key)
                if m is not None:
                    m.write()
