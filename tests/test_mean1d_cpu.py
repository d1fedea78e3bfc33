"""The O(log M) predictive mean of one-input models (csrc/gp_mean1d.h) against the oracle's direct sum, on the
CPU: gpf_selftest_mean1d_host runs the very routine the device kernel calls per query (host instantiation),
so index handling, block walks and cut-offs are checked without a GPU; tests/test_gp_gpu.py then holds the
device kernel to the same oracle."""
import ctypes as C

import numpy as np
import pytest

from gapflow_b200 import _capi
from oracle import gp as G
from tests.gp_cases import query_set, training_set


def mean1d(X, log_amp, log_scale, alpha, xq, grad=True):
    lib = _capi.load()
    X = np.ascontiguousarray(X, dtype=np.float64).ravel()
    a = np.ascontiguousarray(alpha, dtype=np.float64).ravel()
    xq = np.ascontiguousarray(xq, dtype=np.float64).ravel()
    m, g = np.empty(len(xq)), np.empty(len(xq))
    _capi.check(lib.gpf_selftest_mean1d_host(X.ctypes.data, len(X), float(log_amp), float(log_scale), a.ctypes.data,
                                             xq.ctypes.data, len(xq), m.ctypes.data, g.ctypes.data if grad else None))
    return m, g


def reference(X, log_amp, log_scale, alpha, xq):
    ls = np.array([log_scale])
    m = G.repredict_mean(X[:, None], log_amp, ls, alpha[:, None], xq[:, None])[0]
    g = G.dmean_dx0(X[:, None], log_amp, ls, alpha[:, None], xq[:, None])
    return m, g


@pytest.mark.parametrize("M,model,shift", [(150, 'press_wide', 0.0), (2000, 'press', 0.0), (10000, 'press', 0.0),
                                           (2000, 'press', -2.0), (500, 'press_wide', -3.0), (700, 'press_wide', 1.5)])
def test_mean1d_matches_direct_sum_on_trained_models(M, model, shift):
    """alpha of a real factorisation (|alpha| up to ~300 with cancelling signs): mean and slope within 1e-10 of
    the direct O(M) sum, one block to many blocks (length scales shrunk by e^shift)."""
    c = training_set(M, model)
    ls = c['log_scale'] + shift
    L = G.build(c['X'], c['Yerr'], c['log_amp'], ls)
    al = G.alpha(L, c['Y'])[:, 0]
    xq = query_set(3000, c)[:, 0]
    m, g = mean1d(c['X'][:, 0], c['log_amp'], ls[0], al, xq)
    mr, gr = reference(c['X'][:, 0], c['log_amp'], ls[0], al, xq)
    assert np.abs(m - mr).max() <= 1e-10 * np.abs(mr).max()
    assert np.abs(g - gr).max() <= 2e-10 * np.abs(gr).max()


def test_mean1d_edge_cases():
    rng = np.random.default_rng(1)
    # one training point; queries on it, beside it, and so far away that exp(-r) is exactly 0
    X, al = np.array([0.3]), np.array([2.0])
    xq = np.array([0.3, 0.30001, 0.2, 1e4, -1e4])
    m, g = mean1d(X, 0.1, -2.0, al, xq)
    mr, gr = reference(X, 0.1, -2.0, al, xq)
    assert np.allclose(m, mr, rtol=1e-13, atol=0) and np.allclose(g, gr, rtol=1e-12, atol=1e-300)
    assert m[3] == 0.0 and m[4] == 0.0
    # duplicated and unsorted training inputs, queries exactly on training points and outside the range,
    # length scale so small that the points fall into hundreds of blocks and most blocks are cut off
    X = np.concatenate([rng.uniform(0, 1, 300), rng.uniform(0, 1, 50).repeat(2)])
    rng.shuffle(X)
    al = rng.standard_normal(len(X)) * 10.0
    xq = np.concatenate([X[:100], rng.uniform(-0.5, 1.5, 500), [X.min(), X.max(), X.min() - 1e-12, X.max() + 1e-12]])
    for ls in (0.0, -3.0, -7.0):
        m, g = mean1d(X, 0.0, ls, al, xq)
        mr, gr = reference(X, 0.0, ls, al, xq)
        assert np.abs(m - mr).max() <= 1e-12 * np.abs(al).sum(), ls
        assert np.abs(g - gr).max() <= 1e-12 * np.abs(al).sum() * np.exp(-ls), ls
    # the slope output is optional
    m2, _ = mean1d(X, 0.0, -3.0, al, xq, grad=False)
    assert np.array_equal(m2, mean1d(X, 0.0, -3.0, al, xq)[0])


# ---- O(log M) predictive variance of one-input models (csrc/gp_var1d.h) -------------------------------------
def var1d(X, log_amp, log_scale, yerr, xq, noise):
    lib = _capi.load()
    X = np.ascontiguousarray(X, dtype=np.float64).ravel()
    xq = np.ascontiguousarray(xq, dtype=np.float64).ravel()
    v = np.empty(len(xq))
    _capi.check(lib.gpf_selftest_var1d_host(X.ctypes.data, len(X), float(log_amp), float(log_scale), float(yerr),
                                            xq.ctypes.data, len(xq), float(noise), v.ctypes.data))
    return v


@pytest.mark.parametrize("M,model,shift", [(150, 'press_wide', 0.0), (2000, 'press', 0.0), (2000, 'press', -2.0),
                                           (500, 'press_wide', -3.0), (700, 'press_wide', 1.5)])
def test_var1d_matches_dense_solve(M, model, shift):
    """Kalman / two-filter variance == amp + noise - ||L^-1 Ks||^2 (gp.py:509-522) to 1e-12 of the kernel amplitude
    (the dense expression itself cancels amp against the quadratic form), same arg-max."""
    c = training_set(M, model)
    ls = c['log_scale'] + shift
    L = G.build(c['X'], c['Yerr'], c['log_amp'], ls)
    al = G.alpha(L, c['Y'])
    xq = query_set(2000, c)[:, 0]
    _, vref = G.repredict_mean_var(c['X'], c['log_amp'], ls, L, al, 1.5e-8, xq[:, None])
    v = var1d(c['X'][:, 0], c['log_amp'], ls[0], c['Yerr'], xq, 1.5e-8)
    amp = np.exp(c['log_amp'])
    assert np.abs(v - vref).max() <= 1e-12 * amp
    assert np.argmax(v) == np.argmax(vref)
    assert v.min() > 0.0


def test_var1d_edge_cases():
    rng = np.random.default_rng(2)
    X = np.concatenate([rng.uniform(0, 1, 200), rng.uniform(0, 1, 30).repeat(2)])       # duplicated inputs
    rng.shuffle(X)
    xq = np.concatenate([X[:50], rng.uniform(-0.5, 1.5, 300), [X.min(), X.max(), -1e4, 1e4]])
    for ls, yerr, la in ((0.0, 1e-2, 0.0), (-3.0, 1e-3, 0.7), (-6.0, 5e-2, -1.0)):
        Lf = G.build(X[:, None], yerr, la, np.array([ls]))
        _, vref = G.repredict_mean_var(X[:, None], la, np.array([ls]), Lf, np.zeros((len(X), 1)), 0.0, xq[:, None])
        v = var1d(X, la, ls, yerr, xq, 0.0)
        assert np.abs(v - vref).max() <= 1e-11 * np.exp(la), (ls, np.abs(v - vref).max())
        assert v[-1] == np.exp(la) and v[-2] == np.exp(la)          # far from every point: the prior variance
    # one training point
    v = var1d(np.array([0.3]), 0.2, -1.0, 0.1, np.array([0.3, 0.4, -5.0]), 0.0)
    L1 = G.build(np.array([[0.3]]), 0.1, 0.2, np.array([-1.0]))
    _, vref = G.repredict_mean_var(np.array([[0.3]]), 0.2, np.array([-1.0]), L1, np.zeros((1, 1)), 0.0,
                                   np.array([[0.3], [0.4], [-5.0]]))
    assert np.allclose(v, vref, rtol=1e-13, atol=1e-15)
