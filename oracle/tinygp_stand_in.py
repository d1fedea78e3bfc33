"""Stand-ins for tinygp, jaxopt and the few jax functions that the reference's GP-coupled step calls -- TEST
INFRASTRUCTURE ONLY (used by oracle/ref_problem.py to EXECUTE GaPFlow/models/gp.py and stress.py in this container).

This is NOT tinygp: tinygp (>= 0.3.0), jax and jaxopt are not installed here, and GP parity stays "unpinned at the
tinygp boundary" (oracle/gp.py header).  What this file restates is the part of tinygp's published interface that the
reference's call sites use (gp.py:509-603, stress.py:260-268, 522-537, 586-598), in NumPy / SciPy:

    kernels.stationary.Matern32(distance=kernels.distance.L2Distance())   k(r) = (1 + sqrt(3) r) exp(-sqrt(3) r)
    transforms.Linear(scale, kernel)                                      k(scale * x1, scale * x2)
    scalar * kernel                                                       Product(Constant(scalar), kernel)
    GaussianProcess(kernel, X, diag=d)      zero mean, K = k(X, X) + d I, dense lower Cholesky (DirectSolver)
        .X, .kernel(X1, X2) matrix, .kernel(X) diagonal, .solver.solve_triangular(B) = L^-1 B,
        .log_probability(y), .predict(y, X_test) = conditional mean,
        .condition(y, X_test).gp with .loc, .variance (diagonal of the conditioned covariance + noise),
        .mean_function.alpha = K^-1 y, .noise.diag (tinygp's default for diag=None: sqrt(machine epsilon))

so that what EXECUTES around it is the reference's own code: feature matrix, scaling, targets, caching, variance
tolerance, active-learning loop, pauses, scatter of the means into the closure fields.  What cannot be reproduced is made
an explicit input, as in oracle/surrogate.py: the optimiser's result (jaxopt.ScipyMinimize -> ``HOOKS['optimizer']``), the
mock MD's three normal draws (``HOOKS['draws']``), the random initial design (one seeded NumPy generator per
``jax.random.key``).  ``jax.grad`` is a complex-step derivative (exact to rounding for the analytic kernel above),
``jax.vmap`` a Python loop.
"""
import types

import numpy as np
from scipy.linalg import cholesky, cho_solve, solve_triangular

SQRT3 = np.sqrt(3.0)
JITTER = float(np.sqrt(np.finfo(np.float64).eps))
HOOKS = {'optimizer': None, 'draws': (0.0, 0.0, 0.0)}


# ---- kernels --------------------------------------------------------------------------------------------------------
class _Kernel:
    __array_ufunc__ = None                     # so that `numpy_scalar * kernel` reaches __rmul__

    def __call__(self, X1, X2=None):
        if X2 is None:                         # tinygp: the diagonal k(x_n, x_n)
            return np.array([self.matrix(x[None, :], x[None, :])[0, 0] for x in np.atleast_2d(X1)])
        return self.matrix(np.atleast_2d(X1), np.atleast_2d(X2))

    def __rmul__(self, value):
        return Product(Constant(value), self)

    __mul__ = __rmul__


class Constant(_Kernel):
    def __init__(self, value):
        self.value = value

    def matrix(self, X1, X2):
        return self.value * np.ones((len(X1), len(X2)))


class Product(_Kernel):
    def __init__(self, kernel1, kernel2):
        self.kernel1, self.kernel2 = kernel1, kernel2

    def matrix(self, X1, X2):
        return self.kernel1.matrix(X1, X2) * self.kernel2.matrix(X1, X2)


class L2Distance:
    def distance(self, X1, X2):
        return np.sqrt(((X1[:, None, :] - X2[None, :, :])**2).sum(-1))


class Matern32(_Kernel):
    def __init__(self, scale=1.0, distance=None):
        self.scale, self.dist = scale, distance or L2Distance()

    def matrix(self, X1, X2):
        a = SQRT3 * self.dist.distance(X1, X2) / self.scale
        return (1.0 + a) * np.exp(-a)


class Linear(_Kernel):
    def __init__(self, scale, kernel):
        self.scale, self.kernel = scale, kernel

    def matrix(self, X1, X2):
        return self.kernel.matrix(X1 * self.scale, X2 * self.scale)


# ---- the GP object --------------------------------------------------------------------------------------------------
class _Solver:
    def __init__(self, L):
        self.L = L

    def solve_triangular(self, B):
        return solve_triangular(self.L, B, lower=True)


class GaussianProcess:
    def __init__(self, kernel, X, diag=None):
        self.kernel, self.X = kernel, np.asarray(X, dtype=float)
        d = JITTER if diag is None else diag
        K = kernel(self.X, self.X)
        K[np.diag_indices_from(K)] += d
        self.solver = _Solver(cholesky(K, lower=True))
        self.noise = types.SimpleNamespace(diag=np.full(len(self.X), d))

    def _alpha(self, y):
        return cho_solve((self.solver.L, True), np.asarray(y, dtype=float))

    def log_probability(self, y):
        y = np.asarray(y, dtype=float)
        L = self.solver.L
        return -0.5 * y @ self._alpha(y) - np.sum(np.log(np.diag(L))) - 0.5 * len(y) * np.log(2.0 * np.pi)

    def predict(self, y, X_test):
        return self.kernel(self.X, X_test).T @ self._alpha(y)

    def condition(self, y, X_test):
        alpha = self._alpha(y)
        Ks = self.kernel(self.X, X_test)
        A = self.solver.solve_triangular(Ks)
        cov = self.kernel(X_test, X_test) - A.T @ A            # tinygp builds the full conditioned covariance
        cond = types.SimpleNamespace(loc=Ks.T @ alpha, variance=np.diag(cov) + JITTER,
                                     mean_function=types.SimpleNamespace(alpha=alpha),
                                     noise=types.SimpleNamespace(diag=np.full(len(X_test), JITTER)))
        return types.SimpleNamespace(gp=cond)


# ---- jaxopt, jax.grad / vmap, jax.random ----------------------------------------------------------------------------
class ScipyMinimize:
    """jaxopt.ScipyMinimize(fun=...).run(params_init, **kw) -> (params, state.fun_val): the result of the optimiser is an
    explicit input of the comparison (HOOKS['optimizer'](params_init, **kw) -> params), the objective is evaluated."""

    def __init__(self, fun=None, **kw):
        self.fun = fun

    def run(self, params_init, **kw):
        params = HOOKS['optimizer'](params_init, **kw)
        return types.SimpleNamespace(params=params, state=types.SimpleNamespace(fun_val=self.fun(params, **kw)))


def grad(f):
    def df(x, h=1e-30):
        x = np.asarray(x, dtype=float)
        out = np.empty(x.shape)
        for k in range(x.size):
            z = x.astype(complex)
            z[k] += 1j * h
            out[k] = np.imag(f(z)) / h
        return out
    return df


def vmap(f):
    return lambda X: np.stack([f(x) for x in X])


class _Key:
    def __init__(self, seed):
        self.rng = np.random.default_rng(int(seed))


class Random:
    """jax.random as db.py:313-333 and md/mock.py:82-88 use it.  ``current`` is the generator of the latest key(): the
    unseeded Latin hypercube of db.py:313 draws from it too (oracle/ref_problem.py patches db.qmc accordingly)."""
    current = None
    _n = 0

    @classmethod
    def key(cls, seed):
        k = _Key(seed)
        cls.current, cls._n = k.rng, 0
        return k

    @staticmethod
    def split(key, num=2):
        return tuple(key for _ in range(num))

    @classmethod
    def normal(cls, key, shape=()):
        d = HOOKS['draws'][cls._n % 3]
        cls._n += 1
        return d

    @staticmethod
    def uniform(key, shape, minval, maxval):
        return key.rng.uniform(np.broadcast_to(minval, shape), np.broadcast_to(maxval, shape))

    @staticmethod
    def choice(key, n, shape, replace=False):
        return key.rng.choice(n, size=shape[0], replace=replace)
