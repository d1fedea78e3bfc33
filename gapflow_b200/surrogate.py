"""GP-surrogate time step: host orchestration of the reference's ``GaussianProcessSurrogate`` models
around the CUDA library (no CPU path: every array operation on the cells runs in libgapflow_b200.so).

What maps to what (reference file:line):

* ``Mock``            GaPFlow/md/mock.py:33-107   mock MD: stress_bottom / stress_top / eos_pressure at the
                                                 sampled states, evaluated on the device in one batch
                                                 (gpf_stress_bottom, gpf_stress_top, gpf_eos_pressure)
* ``Database``        GaPFlow/db.py:46-369        training arrays, max-abs normalisation, initial design,
                                                 add_data (host: M x 7 and M x 13 arrays)
* ``SurrogateModel``  GaPFlow/models/gp.py:50-506 + stress.py:38-362 (WallStress), :462-622 (Pressure):
                      targets and scales, _train (device Cholesky + device objective/gradient under SciPy
                      BFGS), _infer (gpf_gp_mean / gpf_gp_mean_var on features built on the device),
                      active learning (device arg-max of the variance), pause logic, history
* ``SurrogateDriver`` GaPFlow/problem.py:412-443, 509-586: _pre_run and update with GP closures:
                          features_of(rel) -> gp mean [, variance] -> gpf_fv_set_surrogate -> gpf_fv_pass
                      for both passes, then the record fold with the GP sound speed
                      (gpf_fv_finish_surrogate, stress.py:522-537)

Inputs the reference draws from JAX's PRNG or leaves unseeded are explicit here (``Mock(draws=...)``, the
seeded design in ``Database.initialize``); tinygp's conditioned-GP jitter is ``TINYGP_JITTER``; see
oracle/surrogate.py, which restates the same step in NumPy for the tests.
"""
import ctypes as C

import numpy as np
import torch

from . import _capi
from . import fv as _fv
from .gp import GPModel

TINYGP_JITTER = float(np.sqrt(np.finfo(np.float64).eps))


def _normalizer(x):
    """db.py:264-266: column max-abs floored at 1e-12; host arrays or CUDA tensors (evaluated where the data is)."""
    if isinstance(x, torch.Tensor):
        return torch.clamp(x.abs().amax(dim=0), min=1e-12)
    return np.maximum(np.max(np.abs(x), axis=0), 1e-12)


class Mock:
    """md/mock.py:33-107.  ``draws``: the three standard-normal numbers the reference takes from
    jax.random.key(123) (the same at every call, mock.py:82-88)."""
    is_mock = True

    def __init__(self, prop, geo, gp, draws=(0.0, 0.0, 0.0), device=None):
        self.noise = (gp['press']['obs_stddev'] if gp['press_gp'] else 0.,
                      gp['shear']['obs_stddev'] if gp['shear_gp'] else 0.)
        self.prop, self.geo, self.draws = prop, geo, draws
        self.params = dict(prop)
        self.device = device

    def run_batch(self, X):
        """X (n, 7) -> Y (n, 13), Yerr (n, 13): one device evaluation of the three closures.  A CUDA tensor in gives
        CUDA tensors out (nothing crosses the host: large synthetic training sets); a host array in
        gives host arrays out (what ``Database.add_data`` keeps, db.py:343-369)."""
        dev = self.device or torch.device('cuda', torch.cuda.current_device())
        on_device = isinstance(X, torch.Tensor) and X.is_cuda
        Xd = X.to(torch.float64) if on_device else torch.as_tensor(np.ascontiguousarray(np.atleast_2d(X), dtype=np.float64), device=dev)
        Xd = Xd.reshape(-1, Xd.shape[-1])
        q, h, Ls = Xd[:, :3].T.contiguous(), Xd[:, 3:6].T.contiguous(), Xd[:, 6].contiguous()
        U, V, eta, zeta = self.geo['U'], self.geo['V'], self.prop['shear'], self.prop['bulk']
        n_p, n_s0, n_s1 = (d * s for d, s in zip(self.draws, (self.noise[0], self.noise[1], self.noise[1])))
        bot = _fv.stress_bottom(q, h, U, V, eta, zeta, Ls) + n_s0
        top = _fv.stress_top(q, h, U, V, eta, zeta, Ls) + n_s1
        p = _fv.eos_pressure(q[0].contiguous(), self.prop) + n_p
        Yd = torch.cat([p[None, :], bot, top], dim=0).T.contiguous()
        s = self.noise[1]
        err_row = [self.noise[0], 0., 0., 0., s, s, 0., 0., 0., 0., s, s, 0.]
        if on_device:
            return Yd, torch.tensor(err_row, dtype=torch.float64, device=Xd.device).repeat(Yd.shape[0], 1)
        Y = Yd.cpu().numpy()
        return Y, np.tile(np.array(err_row), (len(Y), 1))

    def run(self, X, tag=None):
        Y, Ye = self.run_batch(np.asarray(X, dtype=float)[None, :])
        return Y[0], Ye[0]


class Database:
    """db.py:46-369 without the dtool persistence (out of scope: storage)."""

    def __init__(self, md, db, num_extra_features=1):
        self._md, self._db = md, db
        self.num_features = 6 + num_extra_features
        self._Xtrain = np.empty((0, self.num_features))
        self._Ytrain = np.empty((0, 13))
        self._Ytrain_err = np.empty((0, 13))
        self.X_scale = np.ones(self.num_features)
        self.Y_scale = np.ones(13)
        self.output_path = None

    has_mock_md = property(lambda self: self._md.is_mock)
    size = property(lambda self: self._Xtrain.shape[0])
    Xtrain = property(lambda self: self._Xtrain / self.X_scale)
    Ytrain = property(lambda self: self._Ytrain / self.Y_scale)
    Ytrain_err = property(lambda self: self._Ytrain_err / self.Y_scale)

    def write(self):
        """db.py:268-276."""
        if self.output_path is not None:
            import os
            np.save(os.path.join(self.output_path, "Xtrain.npy"), self._Xtrain)
            np.save(os.path.join(self.output_path, "Ytrain.npy"), self._Ytrain)
            np.save(os.path.join(self.output_path, "Ytrain_err.npy"), self._Ytrain_err)

    def initialize(self, Xtest, dim=1):
        """db.py:278-341: sampling box around the mean state, gap features of randomly chosen cells."""
        from scipy.stats import qmc
        db = self._db
        Nsample = db['init_size'] - self.size
        if Nsample <= 0:
            return
        if dim == 1:
            flux = np.mean(Xtest[:, 1])
            active = [0, 1]
        else:
            flux = np.hypot(np.mean(Xtest[:, 1]), np.mean(Xtest[:, 2]))
            active = [0, 1, 2]
        rho = np.mean(Xtest[:, 0])
        w = db['init_width']
        lo = np.array([(1.0 - w) * rho, 0.5 * flux, -0.5 * flux])[active]
        hi = np.array([(1.0 + w) * rho, 1.5 * flux, 0.5 * flux])[active]
        rng = np.random.default_rng(int(db['init_seed']))
        if db['init_method'] == 'rand':
            samples = rng.uniform(lo, hi, size=(Nsample, len(active)))
        elif db['init_method'] == 'lhc':
            samples = qmc.scale(qmc.LatinHypercube(d=len(active), seed=rng).random(n=Nsample), lo, hi)
        else:
            m = int(np.log2(Nsample))                     # db.py:418-424: the next power of two, not the previous one
            if 2**m != Nsample:
                m = int(np.ceil(np.log2(Nsample)))
            samples = qmc.scale(qmc.Sobol(d=len(active), seed=rng).random_base2(m=m), lo, hi)
            Nsample = samples.shape[0]
        choice = rng.choice(Xtest.shape[0], size=Nsample, replace=False)
        if len(active) == 2:
            samples = np.hstack([samples, np.zeros((Nsample, 1))])
        self.add_data(np.column_stack([samples, Xtest[choice, 3:]]))

    def add_data(self, Xnew):
        """db.py:343-369; the MD results do not depend on the scales, so the rows are run as one batch."""
        Xnew = np.atleast_2d(np.asarray(Xnew, dtype=float))
        if hasattr(self._md, 'run_batch'):
            Y, Ye = self._md.run_batch(Xnew)
        else:
            res = [self._md.run(X, self.size + k + 1) for k, X in enumerate(Xnew)]
            Y, Ye = np.array([r[0] for r in res]), np.array([r[1] for r in res])
        self._Xtrain = np.vstack([self._Xtrain, Xnew])
        self._Ytrain = np.vstack([self._Ytrain, Y])
        self._Ytrain_err = np.vstack([self._Ytrain_err, Ye])
        self.X_scale = _normalizer(self._Xtrain)
        self.Y_scale = _normalizer(self._Ytrain)
        self.write()


class SurrogateModel:
    """One GP model of the reference bound to an ``FVSolver``: kind 'zz' (pressure, n_out = 1), 'xz' /
    'yz' (lower and upper wall shear, n_out = 2, one kernel; gp.py:546-552)."""

    def __init__(self, kind, solver, database, cfg, optimizer=None, jitter=TINYGP_JITTER, coll=None, row0=0):
        self.kind, self.name, self.solver, self.database = kind, kind, solver, database
        self.coll = coll or LocalColl()       # slabs: maxima, arg-max and the new sample are folded over the ranks
        self.row0 = row0                      # first row of this solver's padded array in the global padded array
        if kind == 'zz':
            self.active_dims = [int(d) for d in cfg['active_dims']]
            self._cols = [0]
        else:
            oi = {'xz': 4, 'yz': 3}[kind]                     # stress.py:92
            self._cols = [oi + 1, oi + 7]
            key, default = ('active_dims_x', [0, 1, 3]) if kind == 'xz' else ('active_dims_y', [0, 2, 3])
            self.active_dims = [int(d) for d in cfg.get(key, default)]
        self.n_out = len(self._cols)
        self.atol, self.rtol = cfg['atol'], cfg['rtol']
        self.max_steps, self.pause_steps = cfg['max_steps'], cfg['pause_steps']
        self.use_active_learning = cfg['active_learning']
        self.optimizer, self.jitter = optimizer, jitter
        self.is_gp_model = True
        self._step, self._pause, self.last_fit_train_size = 0, 0, 0
        self.n_train = [0, 0]
        self.gp = None
        self.mean = None              # normalised predictive mean (n_out, Np), device: what the pass kernels read
        self._var = None              # predictive variance (Nx+2, Ny+2), device, in output units
        self.maximum_variance, self.variance_tol, self.argmax = None, None, None
        self.history = {'step': [], 'database_size': [], 'variance': [], 'obs_stddev': [],
                        'maximum_variance': [], 'variance_tol': []}
        for li in self.active_dims:
            self.history[f'lengthscale_{li}'] = []

    # -- inputs and targets (stress.py:194-262, 541-570) ------------------------------------------
    def Xtest(self, rel=0):
        return _features(self.solver, rel, self.database.X_scale, self.active_dims)

    def _Xtest_row(self, i, rel=0):
        """row i (GLOBAL cell index) of the un-normalised 7-feature matrix (gp.py:223-232, 425); with slabs the
        lowest rank that holds the cell supplies it."""
        loc = int(i) - self.row0 * self.solver.shape[1]
        row = None
        if 0 <= loc < self.solver.shape[0] * self.solver.shape[1]:
            row = _features(self.solver, rel, np.ones(7), list(range(7)))[loc].cpu().numpy()
        return next(r for r in self.coll.allgather(row) if r is not None)

    @property
    def Xtrain(self):
        return self.database.Xtrain[:, self.active_dims]

    @property
    def Yscale(self):
        return float(np.max(self.database.Y_scale[self._cols]))

    @property
    def Ytrain(self):
        return self.database._Ytrain[:self.last_fit_train_size][:, self._cols] / self.Yscale

    @property
    def Yerr(self):
        n = self.last_fit_train_size
        if self.kind == 'zz':
            return float(np.mean(self.database.Ytrain_err[:n, 0]))
        return float(np.mean(self.database._Ytrain_err[:n][:, self._cols] / self.Yscale))

    kernel_variance = property(lambda self: float(np.exp(self.gp.log_amp)))
    kernel_lengthscale = property(lambda self: np.exp(-self.gp.log_scale))
    obs_stddev = property(lambda self: self.Yerr)

    @property
    def trusted(self):
        return self.maximum_variance < self.variance_tol

    @property
    def variance(self):
        return self._var

    # -- training (gp.py:290-335) -------------------------------------------------------------------
    def init(self):
        X = self.Xtrain
        self.params_init = (0.0, np.log(np.std(X, axis=0)))      # stress.py:281-284, 592-595
        self._train()
        self._infer()

    def _train(self, reason=0):
        self.last_fit_train_size = self.database.size
        self.n_train[reason] += 1
        X, Y, yerr = self.Xtrain, self.Ytrain, self.Yerr
        if self.gp is None or self.gp.M_max < len(X):
            self.gp = GPModel(max(64, 2 * len(X)), len(self.active_dims), self.n_out, device=self.solver.device)
        if self.optimizer is None:
            self.gp.train(X, Y, yerr, self.params_init[0], self.params_init[1])
        else:
            la, ls = self.optimizer(self)
            self.gp.factor(X, Y, la, ls, yerr)
        if self._step > 0:
            self.write()

    def write(self):
        """gp.py:247-258."""
        h = self.history
        h['step'].append(self._step)
        h['database_size'].append(self.database.size)
        h['variance'].append(self.kernel_variance)
        h['obs_stddev'].append(self.obs_stddev)
        h['maximum_variance'].append(self.maximum_variance)
        h['variance_tol'].append(self.variance_tol)
        for i, li in enumerate(self.active_dims):
            h[f'lengthscale_{li}'].append(self.kernel_lengthscale[i])

    # -- inference (gp.py:343-414) ---------------------------------------------------------------------
    def _infer(self, compute_var=True, rel=0):
        X = self.Xtest(rel)
        if compute_var:
            self.mean, v = self.gp.repredict_mean_var(X, noise=self.jitter)
            ys = self.Yscale
            self._var = v.reshape(self.solver.shape) * ys**2
            # gp.py:406, 424 over the GLOBAL padded array: slab-internal halo rows repeat the neighbour's cells
            # bit for bit, so the smallest global index among the maxima is the reference's first maximum
            gidx = self.gp.argmax + self.row0 * self.solver.shape[1]
            vmax, self.argmax = fold_first_maximum(self.coll.allgather((self.gp._vmax_norm, gidx)))
            self.maximum_variance = vmax * ys**2
            self.variance_tol = max(self.atol * self.Yerr * ys, self.rtol * ys)**2
        else:
            self.mean = self.gp.repredict_mean(X)
        self.mean_scale = self.Yscale          # gp.py:362: the scale at inference time
        return self.mean, self._var

    def _active_learning(self, rel=0):
        """gp.py:419-430: the cell of the largest variance (first maximum, ghost cells included)."""
        self.database.add_data(self._Xtest_row(self.argmax, rel)[None, :])

    def predict(self, predictor=True, compute_var=True, rel=0):
        """gp.py:435-506."""
        if predictor:
            self._step += 1
            self._pause = max(-1, self._pause - 1)
            if self.last_fit_train_size < self.database.size:
                self._train(reason=0)
        self._infer(compute_var=compute_var and predictor, rel=rel)
        if self.use_active_learning and predictor and self._pause < 0:
            counter = 0
            while not self.trusted and counter < self.max_steps:
                counter += 1
                self._active_learning(rel)
                self._train(reason=1)
                self._infer(compute_var=True, rel=rel)
            if counter == self.max_steps:
                self._pause = self.pause_steps
        return self.mean, self._var

    def mean_field(self):
        """predictive mean in output units, (Nx+2, Ny+2) or (2, Nx+2, Ny+2) (gp.py:362)."""
        m = self.mean.reshape(self.n_out, *self.solver.shape) * self.mean_scale
        return m[0] if self.kind == 'zz' else m

    def v_sound2(self, rel=0):
        """stress.py:522-537: max_i d mean / d x_0 * Yscale / X_scale[0]."""
        _, g = self.gp.repredict_mean(self.Xtest(rel), with_gradient_max=True)
        gs = np.array(self.coll.allgather(g))
        g = float('nan') if np.isnan(gs).any() else float(gs.max())
        return g * self.Yscale / float(self.database.X_scale[0])


def fold_first_maximum(pairs):
    """(value, global index) of the first maximum from every rank's (local maximum, its global index): what
    jnp.max / np.argmax return on the whole array (gp.py:406, 424) -- NaN propagates and wins, ties go to the
    smallest index (cells that several ranks hold carry the same value bit for bit)."""
    v = np.array([a for a, _ in pairs], dtype=float)
    if np.isnan(v).any():
        return float('nan'), min(int(g) for a, g in pairs if a != a)
    vmax = float(v.max())
    return vmax, min(int(g) for a, g in pairs if a == vmax)


class LocalColl:
    """host-level collective of a single domain."""
    rank, world = 0, 1

    def allgather(self, x):
        return [x]


class DistColl:
    """the same over torch.distributed (one process per GPU): small Python objects, rank order."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self._dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def allgather(self, x):
        out = [None] * self.world
        self._dist.all_gather_object(out, x, group=self.group)
        return out


class ThreadColl:
    """the same between threads of one process, one thread per rank (all slabs on one GPU): what the
    single-GPU tests run, with ``exchanger(rank, periodic)`` standing in for dist.HaloExchanger."""

    def __init__(self, world):
        import threading
        self.world = world
        self._barrier = threading.Barrier(world)
        self._slots = [None] * world

    class _Rank:
        def __init__(self, parent, rank):
            self.p, self.rank, self.world = parent, rank, parent.world

        def allgather(self, x):
            p = self.p
            p._slots[self.rank] = x
            p._barrier.wait()
            out = list(p._slots)
            p._barrier.wait()
            return out

    def view(self, rank):
        return ThreadColl._Rank(self, rank)

    def exchanger(self, rank, periodic):
        from .dist import HaloExchanger
        coll = self.view(rank)

        class _Ex(HaloExchanger):
            def exchange(self, send, recv):           # all ranks share the legacy default stream: enqueue order is enough
                sends = coll.allgather(send)
                if self.lo is not None:
                    recv[0].copy_(sends[self.lo][1])
                if self.hi is not None:
                    recv[1].copy_(sends[self.hi][0])
                coll.allgather(None)

            def gather(self, rec, out):
                out.copy_(torch.stack(coll.allgather(rec)).reshape(out.shape))
                coll.allgather(None)
        return _Ex(rank, self.world, periodic)


def _features(solver, rel, X_scale, active_dims):
    lib = _capi.load()
    dims = (C.c_int32 * len(active_dims))(*[int(d) for d in active_dims])
    xs = np.ascontiguousarray(np.asarray(X_scale, dtype=np.float64).reshape(7))
    Np = solver.shape[0] * solver.shape[1]
    X = torch.empty((Np, len(active_dims)), dtype=torch.float64, device=solver.device)
    _capi.check(lib.gpf_gp_features_of(solver.h, int(rel), C.c_void_p(xs.ctypes.data), dims, len(active_dims),
                                       C.c_void_p(X.data_ptr()), solver._stream()))
    return X


class SurrogateDriver:
    """``Problem._pre_run`` / ``Problem.update`` with GP closures (problem.py:412-443, 509-586) on an
    ``FVSolver``.  ``gp``: the sanitised gp dict (io.py:395-427); ``database``: a ``Database``."""

    def __init__(self, solver, gp, database, dim, write_freq=1000, optimizer=None, jitter=TINYGP_JITTER,
                 coll=None, row0=0):
        if database is None:
            raise ValueError("a GP configuration needs a Database (GaPFlow/problem.py:77-152)")
        if database.num_features != 7:
            raise NotImplementedError("the FV solver carries one extra feature plane (problem.py:131)")
        self.solver, self.database, self.dim, self.write_freq = solver, database, dim, write_freq
        self.coll, self.row0 = coll or LocalColl(), row0
        self.lib = _capi.load()
        self.models = {}
        kw = dict(optimizer=optimizer, jitter=jitter, coll=self.coll, row0=row0)
        if gp.get('press_gp'):                                   # problem.py:640-660
            self.models['zz'] = SurrogateModel('zz', solver, database, gp['press'], **kw)
        if gp.get('shear_gp'):
            self.models['xz'] = SurrogateModel('xz', solver, database, gp['shear'], **kw)
            if dim == 2:
                self.models['yz'] = SurrogateModel('yz', solver, database, gp['shear'], **kw)
        self._rec = torch.zeros(5, dtype=torch.float64, device=solver.device)
        self.ready = False

    def _order(self):
        return [self.models[k] for k in ('zz', 'xz', 'yz') if k in self.models]

    def _bind(self):
        """hand the pass kernels the current normalised means (read when the pass runs)."""
        def arg(k):
            m = self.models.get(k)
            return (C.c_void_p(m.mean.data_ptr()), m.mean_scale) if m is not None else (None, 1.0)
        (p, ps), (xz, xs), (yz, ys) = arg('zz'), arg('xz'), arg('yz')
        _capi.check(self.lib.gpf_fv_set_surrogate(self.solver.h, p, ps, xz, xs, yz, ys))

    def _c2(self, rel):
        return self.models['zz'].v_sound2(rel) if 'zz' in self.models else 0.0

    # -- the three device phases a slab driver replaces ---------------------------------------------
    def _global_features(self):
        """the un-normalised feature matrix of every cell of the global padded array (gp.py:223-232)."""
        return _features(self.solver, 0, np.ones(7), list(range(7))).cpu().numpy()

    def _run_pass(self, i):
        s = self.solver
        _capi.check(self.lib.gpf_fv_pass(s.h, i, s._stream()))

    def _finish(self, kind, mode, c2):
        s = self.solver
        _capi.check(self.lib.gpf_fv_reduce_local(s.h, kind, C.c_void_p(self._rec.data_ptr()), s._stream()))
        _capi.check(self.lib.gpf_fv_finish_surrogate(s.h, mode, C.c_void_p(self._rec.data_ptr()), 1, c2, s._stream()))

    # -- problem.py:412-443, 509-586 -------------------------------------------------------------------
    def pre_run(self):
        X = self._global_features()
        for m in self._order():                                   # problem.py:418-420
            self.database.initialize(X, self.dim)
        for m in self._order():                                   # problem.py:422-424
            m.init()
        self._bind()
        self.ready = True
        self._finish(1, _capi.FIN_PRERUN_KEEP_EKIN, self._c2(0))

    def refresh_scalars(self):
        """scalars of the current q with the GP sound speed (Problem.dt_crit / cfl between steps)."""
        self._finish(1, _capi.FIN_REFRESH, self._c2(0))

    def update(self, step):
        """one Problem.update(); ``step`` = Problem.step before it (problem.py:530)."""
        before_output = (step + 1) % self.write_freq == 0
        for i in (0, 1):
            for m in self._order():                               # problem.py:532-540
                m.predict(predictor=i == 0, compute_var=m.use_active_learning or before_output, rel=i)
            self._bind()
            self._run_pass(i)
        self._finish(0, _capi.FIN_STEP, self._c2(2))

    def closures(self, fields):
        """overlay the GP means on the fixed-form closure fields (stress.py:350-356, 619-620); ``fields`` =
        (pressure, wall_xz, wall_yz, bulk) host arrays of FVSolver.closures()."""
        p, wxz, wyz, bulk = fields
        for k, m in self.models.items():
            mf = m.mean_field().cpu().numpy()
            if k == 'zz':
                p = mf
            else:
                f, oi = (wxz, 4) if k == 'xz' else (wyz, 3)
                f[oi], f[oi + 6] = mf[0], mf[1]
        return p, wxz, wyz, bulk


class SlabSurrogateDriver(SurrogateDriver):
    """The same step on one rank's slab (gapflow_b200.dist.SlabSolver): the training set, the factor and the
    database are replicated on every rank (each rank runs the identical mock-MD call and training), the
    query cells are partitioned -- a rank evaluates the GP means of its own padded rows, halo rows
    included, after the halo exchange of the pass's input.  Folded over the ranks: the largest variance
    and its first cell, the new training sample (from the lowest rank that holds the cell), the largest
    slope of the pressure mean, the loop records.  ``coll``: DistColl (one process per GPU) or a
    ThreadColl view (all slabs in one process)."""

    def __init__(self, slab, gp, database, dim, coll, **kw):
        self.slab = slab
        super().__init__(slab.fv, gp, database, dim, coll=coll, row0=slab.row0, **kw)

    def _global_features(self):
        from .dist import owned_rows
        sl, Ny2 = self.slab, self.solver.shape[1]
        a, b = owned_rows(sl.grid_global['Nx'], sl.rank, sl.world)
        X = _features(self.solver, 0, np.ones(7), list(range(7)))
        own = X[(a - sl.row0) * Ny2:(b - sl.row0) * Ny2].cpu().numpy()
        return np.vstack(self.coll.allgather(own))

    def _run_pass(self, i):
        self.slab._pass(i)
        self.slab._exchange(i + 1)            # halo rows of the pass's output: the next pass's (or step's) input

    def _finish(self, kind, mode, c2):
        sl = self.slab
        sl._record(kind)
        sl.ex.gather(sl.rec, sl.recs)
        _capi.check(self.lib.gpf_fv_finish_surrogate(sl.h, mode, C.c_void_p(sl.recs.data_ptr()), sl.world, c2, sl._s()))
