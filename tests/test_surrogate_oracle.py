"""CPU tests of the GP-coupled step's oracle (oracle/surrogate.py) and of the host logic of
gapflow_b200/surrogate.py that needs no device (Database, the model's targets and scales)."""
import numpy as np
import pytest

from oracle import closures as OC
from oracle import gp as OG
from oracle.fv import FVOracle
from oracle import surrogate as OS
from tests import surrogate_cases as SC


def _oracle(name, optimizer=SC.fixed_optimizer, draws=SC.DRAWS, **kw):
    d, _ = SC.inputs(name, **kw)
    g = d['grid']
    extra = SC.slip_length(g['Nx'], g['Ny'])
    md = OS.MockMD(d['properties'], d['geometry'], d['gp'], draws=draws)
    db = OS.Database(md, d['db'])
    o = OS.SurrogateFVOracle(g, d['geometry'], d['properties'], d['numerics'], d['gp'], db, extra=extra,
                             write_freq=d['options']['write_freq'], optimizer=optimizer)
    o.q[...] = SC.perturb(o.q)
    o.fill_ghosts()
    return d, o, db


def test_mock_md_is_the_fixed_form_closures_plus_constant_noise():
    """md/mock.py:81-107: Y = [p, tau_bot (6), tau_top (6)], the SAME three draws at every call."""
    d, _ = SC.inputs('slip1d')
    md = OS.MockMD(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS)
    X = np.array([0.81, 0.05, 0.001, 12., 0.01, 0., 0.4])
    Y, Ye = md.run(X, 1)
    assert Y.shape == (13,) and Ye.shape == (13,)
    g, p = d['geometry'], d['properties']
    sb = OC.stress_bottom(X[:3], X[3:6], g['U'], g['V'], p['shear'], p['bulk'], X[6])
    st = OC.stress_top(X[:3], X[3:6], g['U'], g['V'], p['shear'], p['bulk'], X[6])
    assert Y[0] == OC.eos_pressure(X[0:1], p)[0] + SC.DRAWS[0] * 2e-2
    assert np.array_equal(Y[1:7], sb + SC.DRAWS[1] * 4e-3) and np.array_equal(Y[7:], st + SC.DRAWS[2] * 4e-3)
    assert np.array_equal(Ye, [2e-2, 0, 0, 0, 4e-3, 4e-3, 0, 0, 0, 0, 4e-3, 4e-3, 0])
    Y2, _ = md.run(X, 2)
    assert np.array_equal(Y, Y2)


def test_database_design_and_scales():
    """db.py:264-266, 278-369: init_size rows, sampling box, gap features of existing cells, max-abs scales."""
    d, o, db = _oracle('slip1d', nx=40)
    X = o.models['zz']._Xtest
    db.initialize(X, 1)
    assert db.size == 10
    rho, flux = X[:, 0].mean(), X[:, 1].mean()
    w = d['db']['init_width']
    assert np.all(db._Xtrain[:, 0] >= (1 - w) * rho) and np.all(db._Xtrain[:, 0] <= (1 + w) * rho)
    assert np.all(db._Xtrain[:, 1] >= 0.5 * flux) and np.all(db._Xtrain[:, 1] <= 1.5 * flux)
    assert np.all(db._Xtrain[:, 2] == 0.0)
    rows = {tuple(r) for r in X[:, 3:]}
    assert all(tuple(r) in rows for r in db._Xtrain[:, 3:])
    assert np.array_equal(db.X_scale, np.maximum(np.abs(db._Xtrain).max(0), 1e-12))
    assert np.array_equal(db.Y_scale, np.maximum(np.abs(db._Ytrain).max(0), 1e-12))
    db.initialize(X, 1)                                   # already large enough: no-op
    assert db.size == 10
    db.add_data(X[3][None, :])
    assert db.size == 11 and np.array_equal(db._Xtrain[-1], X[3])


def test_product_database_matches_oracle_database():
    """gapflow_b200.surrogate.Database (host part) == oracle Database for the same MD runner."""
    from gapflow_b200.surrogate import Database
    d, o, db = _oracle('slip2d', nx=12, ny=10)
    X = o.models['zz']._Xtest
    md = OS.MockMD(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS)
    pdb = Database(md, d['db'])
    db.initialize(X, 2)
    pdb.initialize(X, 2)
    assert pdb.size == db.size == 10
    for a, b in ((pdb._Xtrain, db._Xtrain), (pdb._Ytrain, db._Ytrain), (pdb._Ytrain_err, db._Ytrain_err),
                 (pdb.X_scale, db.X_scale), (pdb.Y_scale, db.Y_scale)):
        assert np.array_equal(a, b)
    assert np.all(pdb._Xtrain[:, 2] != 0.0)               # 2D: the y flux is sampled too (db.py:300-302)


@pytest.mark.parametrize('method, size, want', [('rand', 7, 7), ('lhc', 9, 9), ('sobol', 8, 8), ('sobol', 10, 16), ('sobol', 5, 8)])
def test_initial_design_methods_and_sizes(method, size, want):
    """db.py:313-333, 398-425: the three design methods; a Sobol design that is not a power of two is rounded UP to the
    next one (held to the executed reference by tests/test_surrogate_oracle_vs_reference.py, which found a rounding down
    here).  Product and oracle databases draw the same design."""
    from gapflow_b200.surrogate import Database
    d, o, _ = _oracle('slip1d', nx=40)
    cfg = dict(d['db'], init_method=method, init_size=size)
    X = o.models['zz']._Xtest
    md = OS.MockMD(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS)
    db, pdb = OS.Database(md, dict(cfg)), Database(md, dict(cfg))
    db.initialize(X, 1)
    pdb.initialize(X, 1)
    assert db.size == pdb.size == want
    assert np.array_equal(db._Xtrain, pdb._Xtrain) and np.array_equal(db._Ytrain, pdb._Ytrain)


def test_targets_and_scales_of_the_three_models():
    """stress.py:204-262, 551-570."""
    d, o, db = _oracle('slip2d', nx=12, ny=10)
    o.pre_run()
    n = db.size
    zz, xz, yz = o.models['zz'], o.models['xz'], o.models['yz']
    assert zz.Yscale == db.Y_scale[0] and np.array_equal(zz.Ytrain, db._Ytrain[:n, 0] / db.Y_scale[0])
    assert xz.Yscale == max(db.Y_scale[5], db.Y_scale[11]) and yz.Yscale == max(db.Y_scale[4], db.Y_scale[10])
    assert np.array_equal(xz.Ytrain, db._Ytrain[:n][:, [5, 11]] / xz.Yscale)
    assert np.isclose(zz.Yerr, 2e-2 / db.Y_scale[0]) and np.isclose(xz.Yerr, 4e-3 / xz.Yscale)
    assert xz.active_dims == [0, 1, 3, 6] and yz.active_dims == [0, 2, 3, 6] and zz.active_dims == [0, 3]


def test_cached_repredict_equals_conditioning():
    """the reference's own GP test (tests/test_inference.py:86-111): gp.condition == the cached formulas."""
    d, o, db = _oracle('slip1d', nx=30)
    o.pre_run()
    m = o.models['xz']
    Xt = m.Xtest
    mean, var = OG.repredict_mean_var(m.Xtrain, m.log_amp, m.log_scale, m.L, m._alpha(), m.jitter, Xt)
    for k in range(2):
        loc, v = OG.condition(m.Xtrain, m.log_amp, m.log_scale, m.L, m.Ytrain[:, k], m.jitter, Xt)
        assert np.allclose(loc, mean[k], rtol=1e-10, atol=1e-13)
        assert np.allclose(v, var, rtol=1e-7, atol=1e-12)


def test_gp_step_tracks_the_fixed_form_step_when_the_gp_is_well_trained():
    """A GP trained on noise-free mock data of the fixed-form laws reproduces the fixed-form step closely:
    the scatter of stress.py:350-356 / 619-620 puts the means where the laws' values were."""
    d, o, db = _oracle('slip1d', nx=60, atol=1e9, init=40, draws=(0., 0., 0.), optimizer=None)
    for k in ('press', 'shear'):
        d['gp'][k]['active_learning'] = False
    o.models['zz'].use_active_learning = o.models['xz'].use_active_learning = False
    a = FVOracle(d['grid'], d['geometry'], d['properties'], d['numerics'], extra=o.extra)
    a.q[...] = o.q
    o.pre_run()
    assert abs(o.v_sound - a.v_sound) / a.v_sound < 0.05
    a.dt = o.dt
    a.numerics = dict(a.numerics, adaptive=0)
    o.numerics = dict(o.numerics, adaptive=0)
    for _ in range(5):
        o.update()
        a.update()
    err = np.abs(o.q - a.q).max(axis=(1, 2)) / np.maximum(np.abs(a.q).max(axis=(1, 2)), 1e-300)
    assert err[0] < 1e-4 and err[1] < 1e-2, err
    assert np.ptp(o.pressure) > 0 and not np.array_equal(o.pressure, a.pressure)


def test_active_learning_adds_points_retrains_and_pauses():
    """gp.py:458-504: while the largest variance exceeds the tolerance a point is added at its arg-max and
    the model is retrained; the other model retrains (reason DB) at its next predictor call; after
    max_steps misses the model pauses for pause_steps."""
    d, o, db = _oracle('slip1d', nx=40, atol=0.05, al_steps=2)
    o.pre_run()
    n0 = db.size
    xz, zz = o.models['xz'], o.models['zz']
    o.update()
    assert db.size > n0
    assert zz.n_train[1] + xz.n_train[1] == db.size - n0          # one AL training per added point
    assert zz.last_fit_train_size <= db.size and xz.last_fit_train_size == db.size
    added = db._Xtrain[n0:]
    cells = {tuple(np.round(r, 12)) for r in np.vstack([o.extra.reshape(1, -1)]).T}
    assert all((round(x[6], 12),) in cells for x in added)          # slip lengths of real cells
    if not zz.trusted:
        assert zz._pause == zz.pause_steps
    o.update()
    assert zz.last_fit_train_size == db.size                       # caught up with the points xz added
    sizes = [db.size]
    for _ in range(3):
        o.update()
        sizes.append(db.size)
    assert np.isfinite(o.q).all() and o.step == 5
    assert len(zz.history['step']) == zz.n_train[0] + zz.n_train[1] - 1   # every training after init is logged


def test_variance_only_recomputed_before_output_without_active_learning():
    """problem.py:530, stress.py:353-354: compute_var = use_active_learning or one_step_before_output."""
    d, o, db = _oracle('slip1d', nx=30, wf=3)
    for m in o.models.values():
        m.use_active_learning = False
    o.pre_run()
    v0 = o.models['zz']._predictive_var.copy()
    o.update()                       # step 0 -> (0 + 1) % 3 != 0
    assert np.array_equal(o.models['zz']._predictive_var, v0)
    o.update()                       # step 1 -> not yet
    assert np.array_equal(o.models['zz']._predictive_var, v0)
    o.update()                       # step 2 -> (2 + 1) % 3 == 0: recomputed in the predictor pass
    assert not np.array_equal(o.models['zz']._predictive_var, v0)
