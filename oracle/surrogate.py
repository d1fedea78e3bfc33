"""NumPy/SciPy restatement of GaPFlow's GP-coupled time step -- TEST INFRASTRUCTURE ONLY.

Follows (reference file:line):

* GaPFlow/md/mock.py:53-107            Mock MD: stress_bottom / stress_top / eos_pressure at one sample -> MockMD
* GaPFlow/db.py:70-123, 135-170        Database arrays, normalised views                    -> Database
* GaPFlow/db.py:264-266                normaliser (column max-abs floored at 1e-12)
* GaPFlow/db.py:278-341                initialize (sampling box around the mean state)
* GaPFlow/db.py:343-369                add_data (one MD call per row, scales recomputed per row)
* GaPFlow/models/gp.py:290-335         _train                                               -> SurrogateModel._train
* GaPFlow/models/gp.py:343-414         _infer_mean / _infer_mean_var / _infer               -> SurrogateModel._infer
* GaPFlow/models/gp.py:419-430         _active_learning
* GaPFlow/models/gp.py:435-506         predict (retrain on a grown database, active-learning loop, pause)
* GaPFlow/models/stress.py:194-284     WallStress GP targets: columns _out_index+1, _out_index+7, Yscale, Yerr
* GaPFlow/models/stress.py:522-598     Pressure GP targets, v_sound from the GP's slope, init
* GaPFlow/models/stress.py:289-362, 600-622   update: GP means scattered into the closure fields
* GaPFlow/problem.py:412-443, 509-586  _pre_run, update, _post_update with GP models         -> SurrogateFVOracle

PARITY UNPINNED at the JAX boundary (see oracle/gp.py): tinygp, jax and jaxopt are absent.  Everything the reference
itself owns around that boundary IS pinned, by executing it: tests/test_surrogate_oracle_vs_reference.py runs the
reference's Problem / GaussianProcessSurrogate / Pressure / WallStress / Database / Mock classes (oracle/ref_problem.py,
with oracle/tinygp_stand_in.py in tinygp's place) next to this file, step by step.  What the missing packages
leave as INPUTS here, identical for the oracle and the product in every test:
  * the hyperparameters: ``optimizer(model) -> (log_amp, log_scale)``; the default is SciPy BFGS from
    params_init on the restated objective (gp.py:320-321 uses jaxopt.ScipyMinimize);
  * the three standard-normal draws of the mock MD (mock.py:82-88 draws them from jax.random.key(123),
    the same three numbers at every call): ``draws``;
  * the initial design (db.py:313-333 uses an UNSEEDED scipy Latin hypercube and jax.random.choice): a
    seeded NumPy generator;
  * the conditioned GP's diagonal jitter (tinygp's default for ``diag=None``: sqrt(eps), recalled from
    its source): ``TINYGP_JITTER``.
"""
from collections import deque
import numpy as np

from . import closures as C
from . import gp as G
from .fv import FVOracle, predictor_corrector, source

TINYGP_JITTER = float(np.sqrt(np.finfo(np.float64).eps))


class MockMD:
    """md/mock.py:33-107.  ``run(X, tag)`` -> (Y (13,), Yerr (13,))."""
    is_mock = True

    def __init__(self, prop, geo, gp, draws=(0.0, 0.0, 0.0)):
        self.noise = (gp['press']['obs_stddev'] if gp['press_gp'] else 0.,
                      gp['shear']['obs_stddev'] if gp['shear_gp'] else 0.)
        self.prop, self.geo, self.draws = prop, geo, draws

    def run(self, X, tag=None):
        X = np.asarray(X, dtype=float)
        U, V = self.geo['U'], self.geo['V']
        eta, zeta = self.prop['shear'], self.prop['bulk']
        n_p, n_s0, n_s1 = (d * s for d, s in zip(self.draws, (self.noise[0], self.noise[1], self.noise[1])))
        tau_bot = C.stress_bottom(X[:3], X[3:6], U, V, eta, zeta, X[6]) + n_s0
        tau_top = C.stress_top(X[:3], X[3:6], U, V, eta, zeta, X[6]) + n_s1
        press = C.eos_pressure(X[0:1], self.prop) + n_p
        Y = np.hstack([press, tau_bot, tau_top]).T
        s = self.noise[1]
        Ye = np.array([self.noise[0], 0., 0., 0., s, s, 0., 0., 0., 0., s, s, 0.])
        return Y, Ye


class Database:
    """db.py:46-369 without the dtool persistence."""

    def __init__(self, md, db, num_extra_features=1):
        self._md, self._db = md, db
        self.num_features = 6 + num_extra_features
        self._Xtrain = np.empty((0, self.num_features))
        self._Ytrain = np.empty((0, 13))
        self._Ytrain_err = np.empty((0, 13))
        self.X_scale = np.ones(self.num_features)
        self.Y_scale = np.ones(13)

    has_mock_md = property(lambda self: self._md.is_mock)
    size = property(lambda self: self._Xtrain.shape[0])
    Xtrain = property(lambda self: self._Xtrain / self.X_scale)
    Ytrain = property(lambda self: self._Ytrain / self.Y_scale)
    Ytrain_err = property(lambda self: self._Ytrain_err / self.Y_scale)

    def initialize(self, Xtest, dim=1):
        """db.py:278-341 with the seeded NumPy generator named in the module header."""
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
        """db.py:343-369."""
        n = self.size
        for X in np.atleast_2d(Xnew):
            n += 1
            Y, Ye = self._md.run(X, n)
            self._Xtrain = np.vstack([self._Xtrain, X])
            self._Ytrain = np.vstack([self._Ytrain, Y])
            self._Ytrain_err = np.vstack([self._Ytrain_err, Ye])
            self.X_scale = G.normalizer(self._Xtrain)
            self.Y_scale = G.normalizer(self._Ytrain)


def bfgs_optimizer(model):
    """gp.py:320-321 restated with SciPy's BFGS and the analytic gradient, from params_init."""
    la, ls, _ = G.train(model.Xtrain, model.Ytrain, model.Yerr, model.params_init[0], model.params_init[1])
    return la, ls


class SurrogateModel:
    """One GaussianProcessSurrogate (gp.py:50-506) bound to an FV oracle: kind 'zz' (Pressure,
    stress.py:462-622), 'xz' / 'yz' (WallStress, stress.py:38-362)."""

    def __init__(self, kind, fv, database, cfg, optimizer=None, jitter=TINYGP_JITTER):
        self.kind, self.fv, self.database = kind, fv, database
        self.name = kind
        if kind == 'zz':
            self.active_dims = list(cfg['active_dims'])
        else:
            self._out_index = {'xz': 4, 'yz': 3}[kind]
            self.active_dims = list(cfg.get('active_dims_x', [0, 1, 3]) if kind == 'xz' else cfg.get('active_dims_y', [0, 2, 3]))
        self.atol, self.rtol = cfg['atol'], cfg['rtol']
        self.max_steps, self.pause_steps = cfg['max_steps'], cfg['pause_steps']
        self.use_active_learning = cfg['active_learning']
        self.optimizer = optimizer or bfgs_optimizer
        self.jitter = jitter
        self._step, self._pause, self._cache, self.last_fit_train_size = 0, 0, None, 0
        self.n_train = [0, 0]                    # trainings by reason (DB, AL)
        self.history = {'step': [], 'database_size': [], 'variance': [], 'obs_stddev': [],
                        'maximum_variance': [], 'variance_tol': []}
        for li in self.active_dims:
            self.history[f'lengthscale_{li}'] = []

    # -- inputs and targets ---------------------------------------------------------------------
    @property
    def _Xtest(self):                                     # gp.py:223-232
        f = self.fv
        return np.vstack([f.q, f.topo[:3], f.extra]).reshape(self.database.num_features, -1).T

    @property
    def Xtest(self):                                      # stress.py:194-197, 541-544
        return (self._Xtest / self.database.X_scale)[:, self.active_dims]

    @property
    def Xtrain(self):
        return self.database.Xtrain[:, self.active_dims]

    @property
    def _cols(self):
        return [0] if self.kind == 'zz' else [self._out_index + 1, self._out_index + 7]

    @property
    def Yscale(self):                                     # stress.py:236-247, 563-565
        return np.max(self.database.Y_scale[self._cols])

    @property
    def Ytrain(self):                                     # stress.py:204-234, 551-561
        Y = self.database._Ytrain[:self.last_fit_train_size][:, self._cols] / self.Yscale
        return Y[:, 0] if self.kind == 'zz' else Y

    @property
    def Yerr(self):                                       # stress.py:249-262, 567-570
        n = self.last_fit_train_size
        if self.kind == 'zz':
            return np.mean(self.database.Ytrain_err[:n, 0])
        return np.mean(self.database._Ytrain_err[:n][:, self._cols] / self.Yscale)

    @property
    def trusted(self):
        return self.maximum_variance < self.variance_tol

    # -- training and inference -------------------------------------------------------------------
    def init(self):                                       # stress.py:278-287, 586-598
        self.params_init = G.params_init(self.Xtrain)
        self._train()
        self._infer()

    def _train(self, reason=0):                           # gp.py:290-335
        self.last_fit_train_size = self.database.size
        self.n_train[reason] += 1
        self.log_amp, self.log_scale = self.optimizer(self)
        self.log_scale = np.asarray(self.log_scale, dtype=float)
        self.X_fit = self.Xtrain.copy()                   # the GP object keeps ITS inputs (gp.X), scaled as of now
        self.L = G.build(self.X_fit, self.Yerr, self.log_amp, self.log_scale)
        if self._step > 0:
            self.write()
        self._cache = None

    def write(self):                                      # gp.py:247-258
        h = self.history
        h['step'].append(self._step)
        h['database_size'].append(self.database.size)
        h['variance'].append(np.exp(self.log_amp))
        h['obs_stddev'].append(self.Yerr)
        h['maximum_variance'].append(self.maximum_variance)
        h['variance_tol'].append(self.variance_tol)
        for i, li in enumerate(self.active_dims):
            h[f'lengthscale_{li}'].append(np.exp(-self.log_scale[i]))

    def _alpha(self):
        if self._cache is None:                           # gp.py:354-356, 377-379 (condition == re-predict formula)
            self._cache = G.alpha(self.L, self.Ytrain)
        return self._cache

    def _infer(self, compute_var=True):                   # gp.py:388-414
        shape = self.fv.q.shape[-2:]
        a = self._alpha()
        if compute_var:
            m, v = G.repredict_mean_var(self.X_fit, self.log_amp, self.log_scale, self.L, a, self.jitter, self.Xtest)
            self._predictive_var = v.reshape(shape) * self.Yscale**2
            self.maximum_variance = np.max(self._predictive_var)
            self.variance_tol = G.variance_tol(self.atol, self.rtol, self.Yerr, self.Yscale)
        else:
            m = G.repredict_mean(self.X_fit, self.log_amp, self.log_scale, a, self.Xtest)
        mean = m.reshape(-1, *shape) * self.Yscale
        return (mean[0] if self.kind == 'zz' else mean), self._predictive_var

    def _active_learning(self, var):                      # gp.py:419-430
        imax = np.argmax(var)
        self.database.add_data(self._Xtest[imax, :][None, :])

    def predict(self, predictor=True, compute_var=True):  # gp.py:435-506
        if predictor:
            self._step += 1
            self._pause = max(-1, self._pause - 1)
            if self.last_fit_train_size < self.database.size:
                self._train(reason=0)
        m, v = self._infer(compute_var=compute_var and predictor)
        if self.use_active_learning and predictor and self._pause < 0:
            counter = 0
            while not self.trusted and counter < self.max_steps:
                counter += 1
                self._active_learning(v)
                self._train(reason=1)
                m, v = self._infer(compute_var=True)
            if counter == self.max_steps:
                self._pause = self.pause_steps
        return m, v

    def v_sound2(self):                                   # stress.py:522-537
        g = G.dmean_dx0(self.X_fit, self.log_amp, self.log_scale, self._alpha(), self.Xtest)
        return g.max() * self.Yscale / self.database.X_scale[0]


class SurrogateFVOracle(FVOracle):
    """``Problem`` with GP closures (problem.py:77-152, 412-443, 509-586).  ``gp`` is the sanitised gp
    dict (io.py:395-427); ``database`` a ``Database`` around a mock MD."""

    def __init__(self, grid, geo, prop, numerics, gp, database, extra=None, topo=None, write_freq=1000,
                 optimizer=None, jitter=TINYGP_JITTER):
        self.models = {}
        self._ready = False
        super().__init__(grid, geo, prop, numerics, extra=extra, topo=topo, write_freq=write_freq)
        dim = grid.get('dim', 1 if grid['Ny'] == 1 else 2)
        self.dim = dim
        if gp.get('press_gp'):
            self.models['zz'] = SurrogateModel('zz', self, database, gp['press'], optimizer, jitter)
        if gp.get('shear_gp'):                            # problem.py:640-660
            self.models['xz'] = SurrogateModel('xz', self, database, gp['shear'], optimizer, jitter)
            if dim == 2:
                self.models['yz'] = SurrogateModel('yz', self, database, gp['shear'], optimizer, jitter)
        self.database = database
        self.variance = {k: np.zeros(self.q.shape[-2:]) for k in self.models}

    def pre_run(self):
        """problem.py:412-443."""
        for k in ('zz', 'xz', 'yz'):
            if k in self.models:
                self.database.initialize(self.models[k]._Xtest, self.dim)
        for k in ('zz', 'xz', 'yz'):
            if k in self.models:
                self.models[k].init()
        self._ready = True
        self.step, self.simtime, self.residual = 0, 0., 1.
        self.residual_buffer = deque([1.], 5)
        self.dt = self.numerics['CFL'] * self.dt_crit if self.numerics['adaptive'] else self.numerics['dt']

    @property
    def v_sound(self):
        if self._ready and 'zz' in self.models:
            return np.sqrt(self.models['zz'].v_sound2())
        return C.eos_sound_velocity(self.q[0], self.prop).max()

    def closures(self, predictor=False, compute_var=False):
        """stress.py:600-622, 289-362 in the order of problem.py:532-540."""
        super().closures()
        for k, f, idx in (('zz', None, None), ('xz', self.wall_xz, 4), ('yz', self.wall_yz, 3)):
            if k not in self.models:
                continue
            mdl = self.models[k]
            m, v = mdl.predict(predictor=predictor, compute_var=mdl.use_active_learning or compute_var)
            if k == 'zz':
                self.pressure[...] = m
            else:
                f[idx] = m[0]
                f[idx + 6] = m[1]
            self.variance[k][...] = v

    def update(self):
        q0 = self.q.copy()
        dx, dy, dt = self.grid['dx'], self.grid['dy'], self.dt
        before_output = (self.step + 1) % self.write_freq == 0           # problem.py:530
        for i, d in enumerate(self.directions()):
            self.closures(predictor=i == 0, compute_var=before_output)
            fX, fY = predictor_corrector(self.q, self.pressure, self.bulk, d)
            src = source(self.q, self.topo, self.bulk,
                         self.wall_xz[:6] + self.wall_yz[:6], self.wall_xz[6:] + self.wall_yz[6:])
            self.q[...] = self.q - dt * (fX / dx + fY / dy - src)
            self.fill_ghosts()
        self.q[...] = (self.q + q0) / 2.0
        if self.q_is_valid:
            self._post_update()
        else:
            self._finalize(q0)
