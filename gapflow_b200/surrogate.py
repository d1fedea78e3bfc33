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
from collections import deque

import numpy as np
import torch

from . import _capi
from . import fv as _fv
from .gp import GPModel

TINYGP_JITTER = float(np.sqrt(np.finfo(np.float64).eps))


def _normalizer(x):
    """db.py:264-266."""
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
        """X (n, 7) host -> Y (n, 13), Yerr (n, 13) host; one device evaluation of the three closures."""
        dev = self.device or torch.device('cuda', torch.cuda.current_device())
        Xd = torch.as_tensor(np.ascontiguousarray(np.atleast_2d(X), dtype=np.float64), device=dev)
        q, h, Ls = Xd[:, :3].T.contiguous(), Xd[:, 3:6].T.contiguous(), Xd[:, 6].contiguous()
        U, V, eta, zeta = self.geo['U'], self.geo['V'], self.prop['shear'], self.prop['bulk']
        n_p, n_s0, n_s1 = (d * s for d, s in zip(self.draws, (self.noise[0], self.noise[1], self.noise[1])))
        bot = _fv.stress_bottom(q, h, U, V, eta, zeta, Ls) + n_s0
        top = _fv.stress_top(q, h, U, V, eta, zeta, Ls) + n_s1
        p = _fv.eos_pressure(q[0].contiguous(), self.prop) + n_p
        Y = torch.cat([p[None, :], bot, top], dim=0).T.cpu().numpy()
        s = self.noise[1]
        Ye = np.tile(np.array([self.noise[0], 0., 0., 0., s, s, 0., 0., 0., 0., s, s, 0.]), (len(Y), 1))
        return Y, Ye

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
            m = int(np.log2(Nsample))
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

    def __init__(self, kind, solver, database, cfg, optimizer=None, jitter=TINYGP_JITTER):
        self.kind, self.name, self.solver, self.database = kind, kind, solver, database
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
        """row i of the un-normalised 7-feature matrix (gp.py:223-232, 425)."""
        X = _features(self.solver, rel, np.ones(7), list(range(7)))
        return X[int(i)].cpu().numpy()

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
            self.maximum_variance = self.gp._vmax_norm * ys**2
            self.argmax = self.gp.argmax
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
        return g * self.Yscale / float(self.database.X_scale[0])


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

    def __init__(self, solver, gp, database, dim, write_freq=1000, optimizer=None, jitter=TINYGP_JITTER):
        if database is None:
            raise ValueError("a GP configuration needs a Database (GaPFlow/problem.py:77-152)")
        if database.num_features != 7:
            raise NotImplementedError("the FV solver carries one extra feature plane (problem.py:131)")
        self.solver, self.database, self.dim, self.write_freq = solver, database, dim, write_freq
        self.lib = _capi.load()
        self.models = {}
        if gp.get('press_gp'):                                   # problem.py:640-660
            self.models['zz'] = SurrogateModel('zz', solver, database, gp['press'], optimizer, jitter)
        if gp.get('shear_gp'):
            self.models['xz'] = SurrogateModel('xz', solver, database, gp['shear'], optimizer, jitter)
            if dim == 2:
                self.models['yz'] = SurrogateModel('yz', solver, database, gp['shear'], optimizer, jitter)
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

    def pre_run(self):
        s = self.solver
        for m in self._order():                                   # problem.py:418-420
            X = _features(s, 0, np.ones(7), list(range(7))).cpu().numpy()
            self.database.initialize(X, self.dim)
        for m in self._order():                                   # problem.py:422-424
            m.init()
        self._bind()
        self.ready = True
        _capi.check(self.lib.gpf_fv_reduce_local(s.h, 1, C.c_void_p(self._rec.data_ptr()), s._stream()))
        _capi.check(self.lib.gpf_fv_finish_surrogate(s.h, _capi.FIN_PRERUN_KEEP_EKIN, C.c_void_p(self._rec.data_ptr()),
                                                     self._c2(0), s._stream()))

    def refresh_scalars(self):
        """scalars of the current q with the GP sound speed (Problem.dt_crit / cfl between steps)."""
        s = self.solver
        _capi.check(self.lib.gpf_fv_reduce_local(s.h, 1, C.c_void_p(self._rec.data_ptr()), s._stream()))
        _capi.check(self.lib.gpf_fv_finish_surrogate(s.h, _capi.FIN_REFRESH, C.c_void_p(self._rec.data_ptr()),
                                                     self._c2(0), s._stream()))

    def update(self, step):
        """one Problem.update(); ``step`` = Problem.step before it (problem.py:530)."""
        s = self.solver
        before_output = (step + 1) % self.write_freq == 0
        for i in (0, 1):
            for m in self._order():                               # problem.py:532-540
                m.predict(predictor=i == 0, compute_var=m.use_active_learning or before_output, rel=i)
            self._bind()
            _capi.check(self.lib.gpf_fv_pass(s.h, i, s._stream()))
        _capi.check(self.lib.gpf_fv_reduce_local(s.h, 0, C.c_void_p(self._rec.data_ptr()), s._stream()))
        _capi.check(self.lib.gpf_fv_finish_surrogate(s.h, _capi.FIN_STEP, C.c_void_p(self._rec.data_ptr()),
                                                     self._c2(2), s._stream()))

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
