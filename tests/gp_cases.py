"""Frozen synthetic GP training sets.

The reference's LHC sampler is unseeded (db.py:421) and its Mock MD noise uses the JAX PRNG
(md/mock.py:82-88): neither is reproducible here, so the training set is generated once from the
Mock laws (BWR pressure, slip-top wall stresses: md/mock.py:81-107 = the fixed-form closures) with
a seeded NumPy generator, and the same arrays feed the oracle and the CUDA path.  Hyperparameters
are inputs: ``params_init`` of the reference (log_amp = 0, log_scale = log(std(Xtrain)),
stress.py:281-284) unless given."""
import numpy as np

from oracle import closures as C
from oracle import gp as G

PROP = dict(EOS='BWR', T=1.0, gamma=3.0, rho0=0.8, shear=2.15, bulk=0.)
U_WALL, H_GAP = 0.12, 10.0


def training_set(M, model, seed=2025, obs_stddev=None):
    """model: 'press' (active dims [0], n_out 1), 'shear_x' ([0, 1, 6], n_out 2), 'press2' ([0, 3])."""
    rng = np.random.default_rng(seed)
    rho0 = PROP['rho0']
    X = np.zeros((M, 7))
    X[:, 0] = rng.uniform(0.8 * 0.99, 0.8 * 1.01, M) if model != 'press_wide' else rng.uniform(0.6, 0.9, M)
    X[:, 1] = rng.uniform(0.5, 1.5, M) * rho0 * U_WALL / 2
    X[:, 2] = rng.uniform(-0.5, 0.5, M) * rho0 * U_WALL / 2
    X[:, 3] = H_GAP * (1.0 + (0.05 * rng.uniform(-1, 1, M) if model == 'press2' else 0.0))
    X[:, 6] = rng.uniform(0., 1., M)
    q = X[:, :3].T.copy()
    h = X[:, 3:6].T.copy()
    Y = np.zeros((M, 13))
    Y[:, 0] = C.eos_pressure(X[:, 0], PROP)
    sb = C.stress_bottom(q, h, U_WALL, 0., PROP['shear'], PROP['bulk'], X[:, 6])
    st = C.stress_top(q, h, U_WALL, 0., PROP['shear'], PROP['bulk'], X[:, 6])
    Y[:, 1:7] = sb.T
    Y[:, 7:13] = st.T
    if model in ('press', 'press2', 'press_wide'):
        dims, cols = ([0, 3] if model == 'press2' else [0]), [0]
        sd = 2e-2 if obs_stddev is None else obs_stddev
    else:
        dims, cols = [0, 1, 6], [5, 11]            # _out_index + 1, + 7 for xz (stress.py:204-216)
        sd = 4e-3 if obs_stddev is None else obs_stddev
    Y[:, cols] += sd * rng.standard_normal((M, len(cols)))
    X_scale, Y_scale = G.normalizer(X), G.normalizer(Y)
    Yscale = Y_scale[cols].max()
    Xn = (X / X_scale)[:, dims]
    Yn = Y[:, cols] / Yscale
    Yerr = sd / Yscale
    std = np.std(Xn, axis=0)
    log_scale = np.log(np.where(std > 0, std, 1.0))
    return dict(X=Xn, Y=Yn, Yerr=float(Yerr), Yscale=float(Yscale), X_scale=X_scale, dims=dims,
                log_amp=0.0, log_scale=log_scale, n_out=len(cols))


def query_set(N, case, seed=7):
    """Normalised query points covering the training range (plus a little extrapolation)."""
    rng = np.random.default_rng(seed)
    lo, hi = case['X'].min(0), case['X'].max(0)
    span = hi - lo
    return rng.uniform(lo - 0.05 * span, hi + 0.05 * span, (N, len(case['dims'])))
