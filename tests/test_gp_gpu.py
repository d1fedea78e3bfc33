"""GPU parity tests of the GP path (through the C ABI) against the SciPy oracle (oracle/gp.py).

Tolerance (BASELINE.json north_star): GP mean / variance within 1e-9 relative.  "Relative" is taken
against the field's max norm for the mean and against the prior variance amp for the variance
(var = amp + noise - sum v^2 is a cancellation: its own magnitude is not a scale).  The arg-max of the
variance (active learning, gp.py:424) must be the same index.  Parity is unpinned at the tinygp
boundary (see oracle/gp.py); the conditioning identity of the reference's tests/test_inference.py is
checked on the device path too."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

from oracle import gp as G  # noqa: E402
from tests.gp_cases import query_set, training_set  # noqa: E402

REL = 1e-9
NOISE = 1.4901161193847656e-08


def build(case):
    from gapflow_b200.gp import GPModel
    g = GPModel(len(case['X']) + 70, len(case['dims']), case['n_out'])
    g.factor(case['X'], case['Y'], case['log_amp'], case['log_scale'], case['Yerr'])
    return g


@pytest.mark.parametrize("model,M,N", [("press", 37, 1000), ("press", 300, 5000), ("shear_x", 300, 5000),
                                       ("press2", 515, 3001), ("shear_x", 2000, 20000)])
def test_factor_mean_var_vs_oracle(model, M, N):
    c = training_set(M, model)
    g = build(c)
    L = G.build(c['X'], c['Yerr'], c['log_amp'], c['log_scale'])
    a = G.alpha(L, c['Y'])
    Ld, ad = g.get_factor()
    assert np.abs(np.tril(Ld) - L).max() <= REL * np.abs(L).max(), np.abs(np.tril(Ld) - L).max()
    assert np.array_equal(np.triu(Ld, 1), np.zeros_like(Ld))
    # alpha = K^-1 Y is ill-conditioned by itself; what the reference uses it for is the mean
    xt = query_set(N, c)
    m_ref, v_ref = G.repredict_mean_var(c['X'], c['log_amp'], c['log_scale'], L, a, NOISE, xt)
    m, gmax = g.repredict_mean(xt, with_gradient_max=True)
    m = m.cpu().numpy()
    assert m.shape == m_ref.shape
    assert np.abs(m - m_ref).max() <= REL * np.abs(m_ref).max(), np.abs(m - m_ref).max() / np.abs(m_ref).max()
    g_ref = G.dmean_dx0(c['X'], c['log_amp'], c['log_scale'], a, xt)
    assert abs(gmax - g_ref.max()) <= 1e-8 * np.abs(g_ref).max()
    m2, v = g.repredict_mean_var(xt, NOISE)
    v = v.cpu().numpy()
    assert np.array_equal(m2.cpu().numpy(), m)
    amp = np.exp(c['log_amp'])
    assert np.abs(v - v_ref).max() <= REL * amp, np.abs(v - v_ref).max() / amp
    assert g.argmax == int(np.argmax(v_ref))
    assert abs(g._vmax_norm - v_ref.max()) <= REL * amp
    # per-query noise array == scalar noise
    _, v3 = g.repredict_mean_var(xt, torch.full((N,), NOISE, dtype=torch.float64, device='cuda'))
    assert np.array_equal(v3.cpu().numpy(), v)


def test_condition_identity_on_device():
    """tests/test_inference.py:86-111 of the reference: tinygp condition == cached re-predict."""
    c = training_set(60, 'shear_x')
    g = build(c)
    L = G.build(c['X'], c['Yerr'], c['log_amp'], c['log_scale'])
    xt = query_set(400, c)
    m, v = g.repredict_mean_var(xt, NOISE)
    for o in range(2):
        mc, vc = G.condition(c['X'], c['log_amp'], c['log_scale'], L, c['Y'][:, o], NOISE, xt)
        assert np.allclose(m[o].cpu().numpy(), mc, rtol=0, atol=1e-8)
        assert np.allclose(v.cpu().numpy(), vc, rtol=0, atol=1e-8)


def test_refactorisation_after_added_points_and_infer_scaling():
    """Active learning (gp.py:472-504): add the arg-max point, refactor, variance there drops."""
    c = training_set(100, 'press_wide')
    g = build(c)
    xt = query_set(2000, c)
    mean, var = g.infer(xt, (40, 50), c['Yscale'], c['Yerr'], atol=1., rtol=0., noise=NOISE)
    assert mean.shape == (40, 50) and var.shape == (40, 50)
    assert abs(g.maximum_variance - float(var.max())) <= 1e-12 * g.maximum_variance
    assert abs(g.variance_tol - G.variance_tol(1., 0., c['Yerr'], c['Yscale'])) <= 1e-15
    i = g.argmax
    v_before = float(var.reshape(-1)[i]) / c['Yscale']**2
    Xn = np.vstack([c['X'], xt[i:i + 1]])
    Yn = np.vstack([c['Y'], [[0.5]]])
    g.factor(Xn, Yn, c['log_amp'], c['log_scale'], c['Yerr'])
    _, v2 = g.repredict_mean_var(xt, NOISE)
    L = G.build(Xn, c['Yerr'], c['log_amp'], c['log_scale'])
    a = G.alpha(L, Yn)
    _, v_ref = G.repredict_mean_var(Xn, c['log_amp'], c['log_scale'], L, a, NOISE, xt)
    assert np.abs(v2.cpu().numpy() - v_ref).max() <= REL
    assert float(v2[i]) < 0.5 * v_before + 2 * c['Yerr']**2


def test_not_positive_definite_is_reported():
    from gapflow_b200 import _capi
    from gapflow_b200.gp import GPModel
    g = GPModel(10, 1, 1)
    X = np.array([[0.1], [0.1], [0.1]])
    with pytest.raises(_capi.GapflowError):
        g.factor(X, np.ones((3, 1)), 0.0, [0.0], 0.0)     # three identical points, no noise


def test_features_from_solver_storage():
    from gapflow_b200.gp import features
    from tests.test_fv_gpu import make_solver
    s, o = make_solver('slider2d_allDN_slip')      # DH: the FV path is bit-exact, so the features are too
    s.step(2)
    for _ in range(2):
        o.update()
    xs = np.array([880., 45., 9., 2e-5, 0.3, 1., 1e-6])
    for dims in ([0], [0, 1, 6], [0, 2, 6], [0, 3]):
        X = features(s, xs, dims).cpu().numpy()
        assert np.array_equal(X, G.features(o.q, o.topo, o.extra, xs, dims))


@pytest.mark.parametrize("model,M", [("press", 150), ("press2", 257), ("shear_x", 200)])
def test_training_loss_and_gradient_vs_oracle(model, M):
    """Row G9: the objective of _train (gp.py:306-318) and its gradient on the device vs the SciPy
    oracle (1e-9 relative on the loss, 1e-7 of the gradient norm on the gradient)."""
    from gapflow_b200.gp import GPModel
    c = training_set(M, model)
    gp = GPModel(M + 10, len(c['dims']), c['n_out'])
    for la, ls in ((c['log_amp'], c['log_scale']), (0.4, c['log_scale'] - 0.3)):
        gp.factor(c['X'], c['Y'], la, ls, c['Yerr'])
        loss, grad = gp.nlml()
        ref = G.nlml(c['X'], c['Y'], c['Yerr'], la, ls)
        gref = G.nlml_grad(c['X'], c['Y'], c['Yerr'], la, ls)
        assert abs(loss - ref) <= 1e-9 * abs(ref), (loss, ref)
        assert np.abs(grad - gref).max() <= 1e-7 * np.abs(gref).max(), (grad, gref)


def test_train_reaches_the_oracle_optimum():
    """_train (gp.py:290-335): BFGS from params_init on the device objective ends where the same
    optimiser ends on the oracle objective; the final build is the factorisation at the optimum."""
    from gapflow_b200.gp import GPModel
    c = training_set(80, 'press2')
    gp = GPModel(100, len(c['dims']), c['n_out'])
    la, ls, loss = gp.train(c['X'], c['Y'], c['Yerr'])
    la_o, ls_o, loss_o = G.train(c['X'], c['Y'], c['Yerr'])
    assert abs(loss - loss_o) <= 1e-6 * abs(loss_o), (loss, loss_o)
    L, a = gp.get_factor()
    Lr = G.build(c['X'], c['Yerr'], la, ls)
    assert np.abs(L - Lr).max() <= 1e-9 * np.abs(Lr).max()


def test_optin_one_input_variance_kernel_vs_oracle():
    """GPF_GP_VAR1D=1 (read once per process, hence the subprocess): the O(log M) two-filter variance of one-input
    models (csrc/gp_var1d.h) on the device against the oracle's dense solve -- 1e-10 of the amplitude, same arg-max."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, GPF_GP_VAR1D="1")
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "var1d_check.py")], capture_output=True, text=True,
                       timeout=300, env=env, cwd=root)
    assert r.returncode == 0 and "VAR1D_CHECK_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
