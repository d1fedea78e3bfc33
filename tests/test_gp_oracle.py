"""CPU: the GP oracle (oracle/gp.py) against closed forms and the reference's own GP identity.

The reference has no golden GP vectors and tinygp is absent (parity unpinned at that boundary), so
the restatement is pinned by (i) closed-form 1- and 2-point posteriors, (ii) the identity that
tests/test_inference.py:86-111 of the reference asserts -- conditioning formula == cached re-predict
formula -- and (iii) a finite-difference check of the closed-form sound-speed gradient."""
import numpy as np
import pytest

from oracle import gp as G
from tests.gp_cases import query_set, training_set


def test_one_point_posterior_closed_form():
    X = np.array([[0.3]])
    y = np.array([1.7])
    la, ls, yerr = np.log(2.5), np.array([np.log(0.4)]), 0.1
    L = G.build(X, yerr, la, ls)
    a = G.alpha(L, y)
    xt = np.array([[0.3], [0.5], [2.0]])
    k = G.kernel(X, xt, la, ls)[0]
    amp = np.exp(la)
    m, v = G.repredict_mean_var(X, la, ls, L, a, 0.0, xt)
    assert np.allclose(m[0], k * y[0] / (amp + yerr**2), rtol=1e-14)
    assert np.allclose(v, amp - k**2 / (amp + yerr**2), rtol=1e-13)
    r = np.sqrt(3) * np.abs(xt[:, 0] - 0.3) / 0.4
    assert np.allclose(k, amp * (1 + r) * np.exp(-r), rtol=1e-15)


def test_two_point_posterior_closed_form():
    X = np.array([[0.0, 0.0], [1.0, 0.5]])
    Y = np.array([[1.0, -1.0], [2.0, 0.5]])
    la, ls, yerr = 0.3, np.log(np.array([0.7, 1.3])), 0.05
    K = G.kernel(X, X, la, ls) + yerr**2 * np.eye(2)
    L = G.build(X, yerr, la, ls)
    assert np.allclose(L @ L.T, K, rtol=1e-15)
    a = G.alpha(L, Y)
    assert np.allclose(a, np.linalg.inv(K) @ Y, rtol=1e-13)
    xt = np.array([[0.2, 0.1], [0.9, 0.6]])
    ks = G.kernel(X, xt, la, ls)
    m, v = G.repredict_mean_var(X, la, ls, L, a, 1e-8, xt)
    assert np.allclose(m, (ks.T @ np.linalg.inv(K) @ Y).T, rtol=1e-12)
    assert np.allclose(v, np.exp(la) + 1e-8 - np.einsum('ij,ik,kj->j', ks, np.linalg.inv(K), ks), rtol=1e-11)


def test_condition_equals_repredict_like_reference_test_inference():
    for model in ('press', 'shear_x'):
        c = training_set(40, model)
        L = G.build(c['X'], c['Yerr'], c['log_amp'], c['log_scale'])
        a = G.alpha(L, c['Y'])
        xt = query_set(300, c)
        noise = 1.4901161193847656e-08
        m, v = G.repredict_mean_var(c['X'], c['log_amp'], c['log_scale'], L, a, noise, xt)
        for o in range(c['n_out']):
            mc, vc = G.condition(c['X'], c['log_amp'], c['log_scale'], L, c['Y'][:, o], noise, xt)
            assert np.allclose(mc, m[o], rtol=0, atol=1e-8)        # jnp.isclose(..., 0.) of the reference
            assert np.allclose(vc, v, rtol=0, atol=1e-8)
        assert np.allclose(G.repredict_mean(c['X'], c['log_amp'], c['log_scale'], a, xt), m, rtol=1e-15)


def test_sound_speed_gradient_matches_finite_differences():
    c = training_set(30, 'press2')
    L = G.build(c['X'], c['Yerr'], c['log_amp'], c['log_scale'])
    a = G.alpha(L, c['Y'])
    xt = query_set(50, c)
    g = G.dmean_dx0(c['X'], c['log_amp'], c['log_scale'], a, xt)
    eps = 1e-6
    xp, xm = xt.copy(), xt.copy()
    xp[:, 0] += eps
    xm[:, 0] -= eps
    fd = (G.repredict_mean(c['X'], c['log_amp'], c['log_scale'], a, xp)[0]
          - G.repredict_mean(c['X'], c['log_amp'], c['log_scale'], a, xm)[0]) / (2 * eps)
    assert np.allclose(g, fd, rtol=1e-6, atol=1e-7 * np.abs(fd).max())


def test_features_layout():
    rng = np.random.default_rng(0)
    q, topo, extra = rng.random((3, 4, 5)), rng.random((4, 4, 5)), rng.random((1, 4, 5))
    xs = np.arange(1., 8.)
    F = G.features(q, topo, extra, xs, [0, 2, 6])
    assert F.shape == (20, 3)
    assert F[1 * 5 + 3, 0] == q[0, 1, 3] / 1. and F[7, 1] == q[2, 1, 2] / 3. and F[19, 2] == extra[0, 3, 4] / 7.


def test_training_loss_gradient_matches_finite_differences():
    """gp.py:306-318: the oracle's analytic gradient of -sum_o log N(Y_o; 0, K) vs central differences,
    single and two outputs."""
    for model in ('press2', 'shear_x'):
        c = training_set(40, model)
        th = np.concatenate([[0.3], c['log_scale'] + 0.2])
        g = G.nlml_grad(c['X'], c['Y'], c['Yerr'], th[0], th[1:])
        for k in range(len(th)):
            e = np.zeros_like(th)
            e[k] = 1e-6
            fd = (G.nlml(c['X'], c['Y'], c['Yerr'], *(lambda t: (t[0], t[1:]))(th + e))
                  - G.nlml(c['X'], c['Y'], c['Yerr'], *(lambda t: (t[0], t[1:]))(th - e))) / 2e-6
            assert abs(g[k] - fd) <= 1e-5 * max(1.0, abs(fd)), (model, k, g[k], fd)


def test_training_loss_equals_gaussian_log_density():
    """-log N(y; 0, K) written out with the dense inverse and determinant (small M)."""
    c = training_set(12, 'press')
    K = G.kernel(c['X'], c['X'], c['log_amp'], c['log_scale']) + c['Yerr']**2 * np.eye(12)
    y = c['Y'][:, 0]
    ref = 0.5 * y @ np.linalg.solve(K, y) + 0.5 * np.linalg.slogdet(K)[1] + 6 * np.log(2 * np.pi)
    assert abs(G.nlml(c['X'], c['Y'], c['Yerr'], c['log_amp'], c['log_scale']) - ref) <= 1e-10 * abs(ref)


@pytest.mark.parametrize('model, M', [('press', 60), ('press2', 150), ('shear_x', 200)])
def test_oracle_against_an_independent_implementation(model, M):
    """tinygp is absent here (parity unpinned at that boundary, oracle/gp.py header), so the restated algorithm is held
    to an INDEPENDENT implementation of the same published one (Rasmussen & Williams, Alg. 2.1): scikit-learn's
    GaussianProcessRegressor with amp * Matern(nu = 1.5, one length scale per input) and alpha = yerr^2 -- posterior mean,
    predictive variance (without the conditioned GP's jitter), log marginal likelihood and its gradient in the
    log-parameters.  log_scale of gp.py:598-601 is the log of the LENGTH scale (inputs are multiplied by exp(-log_scale))."""
    from sklearn.gaussian_process import GaussianProcessRegressor
    from sklearn.gaussian_process.kernels import ConstantKernel, Matern
    c = training_set(M, model)
    rng = np.random.default_rng(11)
    log_amp = 0.3
    log_scale = c['log_scale'] + 0.2 * rng.standard_normal(len(c['dims']))
    Xq = query_set(400, c)
    kern = ConstantKernel(np.exp(log_amp)) * Matern(length_scale=np.exp(log_scale), nu=1.5)
    L = G.build(c['X'], c['Yerr'], log_amp, log_scale)
    assert np.allclose(G.kernel(c['X'][:50], Xq[:70], log_amp, log_scale), kern(c['X'][:50], Xq[:70]), rtol=1e-12, atol=0)
    loss, grad = 0.0, 0.0
    for o in range(c['n_out']):
        y = c['Y'][:, o]
        ref = GaussianProcessRegressor(kernel=kern, alpha=c['Yerr']**2, optimizer=None, normalize_y=False).fit(c['X'], y)
        m_ref, s_ref = ref.predict(Xq, return_std=True)
        a = G.alpha(L, y)
        mean, var = G.repredict_mean_var(c['X'], log_amp, log_scale, L, a, 0.0, Xq)
        assert np.allclose(mean[0], m_ref, rtol=1e-9, atol=1e-9 * np.abs(m_ref).max()), (model, o)
        # amp - |L^-1 k|^2 cancels near training points: absolute in units of the prior variance
        assert np.abs(var - s_ref**2).max() <= 1e-9 * np.exp(log_amp), (model, o)
        lml, dlml = ref.log_marginal_likelihood(ref.kernel_.theta, eval_gradient=True)
        loss, grad = loss - lml, grad - dlml
    assert np.isclose(G.nlml(c['X'], c['Y'], c['Yerr'], log_amp, log_scale), loss, rtol=1e-10)
    g = G.nlml_grad(c['X'], c['Y'], c['Yerr'], log_amp, log_scale)
    assert np.allclose(g, grad, rtol=1e-7, atol=1e-7 * np.abs(grad).max()), (g, grad)


@pytest.mark.parametrize('model, M', [('press', 60), ('press2', 120), ('shear_x', 150)])
def test_trained_optimum_equals_an_independent_optimiser_s(model, M):
    """gp.py:320-323 minimises the summed negative log likelihood from params_init; scikit-learn maximises the same sum
    (one kernel for all outputs) with L-BFGS-B from the same start.  Same optimum: the objective to 1e-7 relative (a
    length scale that the data do not determine, press2's gap height, may sit anywhere on its plateau)."""
    from sklearn.gaussian_process import GaussianProcessRegressor
    from sklearn.gaussian_process.kernels import ConstantKernel, Matern
    c = training_set(M, model)
    la, ls, loss = G.train(c['X'], c['Y'], c['Yerr'])
    a0, s0 = G.params_init(c['X'])
    kern = ConstantKernel(np.exp(a0), (1e-8, 1e8)) * Matern(length_scale=np.exp(s0), length_scale_bounds=(1e-8, 1e8), nu=1.5)
    ref = GaussianProcessRegressor(kernel=kern, alpha=c['Yerr']**2).fit(c['X'], c['Y'] if c['n_out'] > 1 else c['Y'][:, 0])
    assert np.isclose(loss, -ref.log_marginal_likelihood_value_, rtol=1e-7), (loss, ref.log_marginal_likelihood_value_)
    th = ref.kernel_.theta
    assert np.isclose(G.nlml(c['X'], c['Y'], c['Yerr'], th[0], th[1:]), loss, rtol=1e-7)
    if model != 'press2':
        assert np.allclose(np.concatenate([[la], ls]), th, atol=2e-3), (la, ls, th)
