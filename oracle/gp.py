"""NumPy/SciPy FP64 restatement of GaPFlow's GP surrogate inference -- TEST INFRASTRUCTURE ONLY.

Follows (reference file:line):

* GaPFlow/models/gp.py:576-603   multi_in_single_out: kernel = exp(log_amp) * Linear(exp(-log_scale),
                                 Matern32(L2Distance)), GaussianProcess(kernel, X, diag=yerr**2)
* GaPFlow/models/gp.py:525-535   _repredict_mean:      Ks = k(Xtrain, Xtest); mean = Ks.T @ alpha
* GaPFlow/models/gp.py:509-522   _repredict_mean_var:  v = L^-1 Ks; var = k(x,x) + noise - sum(v**2, 0)
* GaPFlow/models/gp.py:538-570   _predict_*_output:    alpha = K^-1 y per output column (tinygp condition)
* GaPFlow/models/gp.py:406-410   maximum_variance, variance_tol
* GaPFlow/models/gp.py:419-430   _active_learning: argmax of the variance
* GaPFlow/models/stress.py:194-197, 541-544  Xtest = (_Xtest / X_scale)[:, active_dims]
* GaPFlow/models/gp.py:223-232   _Xtest = vstack([solution, topo[:3], extra]).reshape(n_feat, -1).T
* GaPFlow/models/stress.py:522-537  v_sound^2 = max_i d mean / d x_0 * Yscale / X_scale[0]
* GaPFlow/db.py:264-266          normaliser: max(|x|) per column, floored at 1e-12
* GaPFlow/models/gp.py:290-335   _train: loss = -sum_outputs gp.log_probability(Y_o), minimised by
                                 jaxopt.ScipyMinimize (SciPy's default: BFGS) from params_init
* GaPFlow/models/stress.py:281-284, 592-595  params_init = {log_amp: 0, log_scale: log(std(Xtrain, axis=0))}

PARITY UNPINNED at the tinygp boundary: tinygp / JAX / jaxopt are neither installed nor vendored
(pyproject.toml:33-37 pins tinygp>=0.3.0, jax<=0.9.0, jaxopt==0.8.5) and the reference holds no
golden vector for GP values; its only GP test (tests/test_inference.py:86-111) is the identity
"conditioning formula == cached re-predict formula", which ``condition`` below reproduces.  tinygp's
published algorithm as restated here: DirectSolver = dense Cholesky of K + diag; Matern32(r) =
(1 + sqrt(3) r) exp(-sqrt(3) r); Linear(scale, k)(x1, x2) = k(scale*x1, scale*x2); L2Distance =
sqrt(sum((x1 - x2)**2)).  The conditioned GP's diagonal jitter (``noise``) is an INPUT here.
What stands in for the missing pin (next to tests/test_surrogate_oracle_vs_reference.py, which executes the reference's
own code around the tinygp calls): tests/test_gp_oracle.py holds this file to an independent implementation of the same
published algorithm (scikit-learn's GaussianProcessRegressor, amp * Matern(nu=1.5) with one length scale per input,
alpha = yerr^2): mean 1e-9, variance 1e-9 amp, log marginal likelihood 1e-10, its gradient 1e-7.
"""
import numpy as np
from scipy.linalg import cholesky, cho_solve, solve_triangular

SQRT3 = np.sqrt(3.0)


def kernel(X1, X2, log_amp, log_scale):
    """k(X1, X2) -> (len(X1), len(X2)); gp.py:598-601."""
    s = np.exp(-np.asarray(log_scale, dtype=float))
    A = np.asarray(X1, dtype=float) * s
    B = np.asarray(X2, dtype=float) * s
    d2 = ((A[:, None, :] - B[None, :, :])**2).sum(-1)
    a = SQRT3 * np.sqrt(d2)
    return np.exp(log_amp) * (1.0 + a) * np.exp(-a)


def kernel_blocked(X1, X2, log_amp, log_scale, block=2048):
    """Same as ``kernel`` without the (n1, n2, D) temporary."""
    out = np.empty((len(X1), len(X2)))
    for i in range(0, len(X1), block):
        out[i:i + block] = kernel(X1[i:i + block], X2, log_amp, log_scale)
    return out


def build(Xtrain, yerr, log_amp, log_scale):
    """GaussianProcess(kernel, X, diag=yerr**2): returns the lower Cholesky factor of K + yerr^2 I."""
    K = kernel_blocked(Xtrain, Xtrain, log_amp, log_scale)
    K[np.diag_indices_from(K)] += yerr**2
    return cholesky(K, lower=True)


def alpha(L, Y):
    """K^-1 Y per output column (gp.py:546-552); Y (M,) or (M, n_out) -> (M, n_out)."""
    Y = np.asarray(Y, dtype=float)
    if Y.ndim == 1:
        Y = Y[:, None]
    return cho_solve((L, True), Y)


def repredict_mean(Xtrain, log_amp, log_scale, alpha_, Xtest):
    """gp.py:525-535 -> (n_out, N)."""
    Ks = kernel_blocked(Xtrain, Xtest, log_amp, log_scale)
    return (Ks.T @ alpha_).T


def repredict_mean_var(Xtrain, log_amp, log_scale, L, alpha_, noise, Xtest):
    """gp.py:509-522 -> (n_out, N), (N,)."""
    Ks = kernel_blocked(Xtrain, Xtest, log_amp, log_scale)
    mean = Ks.T @ alpha_
    v = solve_triangular(L, Ks, lower=True)
    Kss = np.exp(log_amp) + noise            # k(x, x) = amp for a stationary kernel
    var = Kss - np.sum(v**2, axis=0)
    return mean.T, var


def condition(Xtrain, log_amp, log_scale, L, Y, noise, Xtest):
    """tinygp ``gp.condition(y, Xtest)``: loc and the DIAGONAL of Kss - A^T A + noise, A = L^-1 Ks,
    computed from the full N x N covariance like tinygp does (small N only)."""
    Ks = kernel(Xtrain, Xtest, log_amp, log_scale)
    a = cho_solve((L, True), np.asarray(Y, dtype=float))
    A = solve_triangular(L, Ks, lower=True)
    cov = kernel(Xtest, Xtest, log_amp, log_scale) - A.T @ A
    return Ks.T @ a, np.diag(cov) + noise


def dmean_dx0(Xtrain, log_amp, log_scale, alpha_, Xtest, block=2048):
    """d mean_0 / d x_0 at every test point: what vmap(grad(lambda x: gp.predict(Y, x[None])))(Xtest)[:, 0]
    evaluates (stress.py:522-537), in closed form:
    dk/dx_0 = -3 amp s_0^2 (x_0 - x'_0) exp(-sqrt(3) r)."""
    s = np.exp(-np.asarray(log_scale, dtype=float))
    amp = np.exp(log_amp)
    out = np.empty(len(Xtest))
    A = np.asarray(Xtrain) * s
    for i in range(0, len(Xtest), block):
        B = np.asarray(Xtest[i:i + block]) * s
        diff = B[:, None, :] - A[None, :, :]
        r = np.sqrt((diff**2).sum(-1))
        dk = -3.0 * amp * s[0] * diff[:, :, 0] * np.exp(-SQRT3 * r)      # s_0 * (s_0 x_0 - s_0 x'_0)
        out[i:i + block] = dk @ alpha_[:, 0]
    return out


def features(q, topo, extra, X_scale, active_dims):
    """gp.py:223-232 + stress.py:194-197: (N, D) normalised, column-selected test inputs; row index
    i*(Ny+2)+j, ghost cells included."""
    extra = np.asarray(extra).reshape(-1, *q.shape[1:])
    F = np.vstack([q, topo[:3], extra]).reshape(3 + 3 + extra.shape[0], -1).T
    return (F / np.asarray(X_scale))[:, list(active_dims)]


def normalizer(x):
    """db.py:264-266."""
    return np.maximum(np.max(np.abs(x), axis=0), 1e-12)


def variance_tol(atol, rtol, Yerr, Yscale):
    """gp.py:408-410."""
    return np.maximum(atol * Yerr * Yscale, rtol * Yscale)**2


def nlml(Xtrain, Y, yerr, log_amp, log_scale):
    """gp.py:306-318: -sum_o log N(Y_o; 0, K + yerr^2 I) (tinygp log_probability: -1/2 y^T K^-1 y -
    sum log L_ii - M/2 log 2 pi)."""
    Y = np.asarray(Y, dtype=float)
    if Y.ndim == 1:
        Y = Y[:, None]
    L = build(Xtrain, yerr, log_amp, log_scale)
    a = cho_solve((L, True), Y)
    M = len(Y)
    return 0.5 * np.sum(Y * a) + Y.shape[1] * (np.sum(np.log(np.diag(L))) + 0.5 * M * np.log(2.0 * np.pi))


def nlml_grad(Xtrain, Y, yerr, log_amp, log_scale):
    """Analytic gradient of ``nlml`` w.r.t. (log_amp, log_scale[0..D-1]) -- what JAX autodiff gives the
    reference's optimiser: 1/2 sum_ij (n_out K^-1 - sum_o a_o a_o^T)_ij dK_ij/dtheta."""
    Y = np.asarray(Y, dtype=float)
    if Y.ndim == 1:
        Y = Y[:, None]
    s = np.exp(-np.asarray(log_scale, dtype=float))
    A = np.asarray(Xtrain, dtype=float) * s
    L = build(Xtrain, yerr, log_amp, log_scale)
    a = cho_solve((L, True), Y)
    W = Y.shape[1] * cho_solve((L, True), np.eye(len(Y))) - a @ a.T
    diff2 = (A[:, None, :] - A[None, :, :])**2
    r = np.sqrt(diff2.sum(-1))
    amp = np.exp(log_amp)
    ex = np.exp(-SQRT3 * r)
    g = [0.5 * np.sum(W * amp * (1.0 + SQRT3 * r) * ex)]
    for d in range(A.shape[1]):
        g.append(0.5 * np.sum(W * 3.0 * amp * ex * diff2[:, :, d]))
    return np.array(g)


def params_init(Xtrain):
    """stress.py:281-284."""
    return 0.0, np.log(np.std(np.asarray(Xtrain, dtype=float), axis=0))


def train(Xtrain, Y, yerr, log_amp0=None, log_scale0=None):
    """gp.py:320-323 with SciPy's BFGS and the analytic gradient -> (log_amp, log_scale, loss)."""
    from scipy.optimize import minimize
    D = np.asarray(Xtrain).shape[1]
    a0, s0 = params_init(Xtrain)
    x0 = np.concatenate([[a0 if log_amp0 is None else log_amp0], s0 if log_scale0 is None else log_scale0])
    res = minimize(lambda t: (nlml(Xtrain, Y, yerr, t[0], t[1:]), nlml_grad(Xtrain, Y, yerr, t[0], t[1:])), x0, jac=True,
                   method='BFGS')
    return res.x[0], res.x[1:1 + D], res.fun
