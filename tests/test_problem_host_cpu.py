"""CPU: the HOST logic of the product's ``gapflow_b200.problem.Problem`` -- host edits of ``q`` / ``dt`` /
``kinetic_energy_old``, the stale closure fields, the run loop and its device batches, when frames and history rows are
written and what they hold -- with the device solver replaced by an oracle-backed double (tests/host_double.py), held to
the reference's own ``Problem`` EXECUTED next to it (oracle/ref_problem.py).  The CUDA solver behind the same class is
what tests/test_problem_gpu.py runs; nothing in the product imports the double."""
import contextlib
import csv
import io
import os

import numpy as np
import pytest

from oracle import ref_loader
from tests import cases
from tests.host_double import OracleFVSolver
from tests.test_problem_gpu import DECAY, JOURNAL2D, SOMMERFELD

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")


@pytest.fixture(scope="module")
def ref():
    from oracle import ref_problem
    yield ref_problem.reference_package()
    ref_problem.release()


@pytest.fixture
def Host(monkeypatch):
    import gapflow_b200.problem as P
    monkeypatch.setattr(P, 'FVSolver', OracleFVSolver)
    return P.Problem


def _ref(ref, text, silent=True, **opts):
    with io.StringIO(text) as f, contextlib.redirect_stdout(io.StringIO()):
        rd = ref.read_yaml_input(f)
        rd['options']['silent'] = silent
        rd['options'].update(opts)
        return ref.Problem._from_dict(rd)


def _same(a, r, tag):
    assert np.array_equal(a.q, r.q), tag
    assert a.dt == r.dt and a.step == r.step and a.simtime == r.simtime and a.residual == r.residual, tag
    assert list(a.residual_buffer) == list(r.residual_buffer) and bool(a.converged) == bool(r.converged), tag
    assert a.mass == r.mass and a.kinetic_energy == r.kinetic_energy and a.v_max == r.v_max, tag
    assert a.dt_crit == r.dt_crit and a.cfl == r.cfl and a.pressure.v_sound == r.pressure.v_sound, tag
    assert a.kinetic_energy_old == r.kinetic_energy_old, tag
    # the closure fields as the reference holds them between steps: one pass stale (problem.py:532-540 vs :563)
    assert np.array_equal(a.pressure.pressure, r.pressure.pressure), tag
    assert np.array_equal(a.wall_stress_xz.full, r.wall_stress_xz.full), tag
    assert np.array_equal(a.wall_stress_xz.lower, r.wall_stress_xz.lower), tag
    assert np.array_equal(a.wall_stress_yz.upper, r.wall_stress_yz.upper), tag
    assert np.array_equal(a.bulk_stress.stress, r.bulk_stress.stress), tag
    assert np.array_equal(a.topo.h, r.topo.h) and np.array_equal(a.topo.x, r.topo.x), tag


@pytest.mark.parametrize('name', ['journal2d_periodic', 'slider2d_xDN', 'asym_xy_mc0', 'journal2d_flip_mc0', 'slider2d_barus_eyring'])
def test_update_with_host_edits_like_the_reference(ref, Host, name):
    text = cases.STEP_CASES[name]['yaml']
    a, r = Host.from_string(text), _ref(ref, text)
    bump = 1 + 1e-6 * np.random.default_rng(3).standard_normal(a.q[0].shape)
    for p in (a, r):
        p.q[0] *= bump                              # edit before the first step (tests/test_wave_decay.py:101)
        with contextlib.redirect_stdout(io.StringIO()):
            p._pre_run()
    _same(a, r, 'pre_run')
    for k in range(3):
        a.update()
        r.update()
        _same(a, r, k)
    for p in (a, r):
        p.q[1] += 1e-4                              # edit mid-run, no _pre_run
        p.kinetic_energy_old = p.kinetic_energy     # tests/test_wave_decay.py:102
        p.dt = 0.5 * p.dt
    assert a.dt == r.dt
    for k in range(2):
        a.update()
        r.update()
        _same(a, r, ('edited', k))
    assert a.q_is_valid and not a.q_has_nan and not a.q_has_negative_density
    for p in (a, r):                                # garbage in the ghost cells, then the reference's refill
        p.q[:, 0, :] = 7.
        p.q[:, :, -1] = -3.
        p._communicate_ghost_buffers()
    assert np.array_equal(a.q, r.q)
    a.update()
    r.update()
    _same(a, r, 'refilled')


def test_invalid_step_restores_the_field_and_stops(ref, Host):
    text = cases.STEP_CASES['slider2d_xDN']['yaml']
    a, r = Host.from_string(text), _ref(ref, text)
    for p in (a, r):
        with contextlib.redirect_stdout(io.StringIO()):
            p._pre_run()
            p.update()
            p.dt = 1e-3
            for _ in range(3):
                p.update()
    assert a._stop and r._stop and a.step == r.step
    assert np.array_equal(a.q, r.q) and np.array_equal(a.pressure.pressure, r.pressure.pressure)


@pytest.mark.parametrize('tol, max_it, wf', [(1e-4, 10_000, 1000), (1e-12, 130, 50), (1e-4, 10_000, 37)])
def test_run_loop_and_outputs(ref, Host, tmp_path, tol, max_it, wf):
    """run(): stops where the reference stops (convergence or max_it), whatever the batch length; frames of sol.nc and
    rows of history.csv at the reference's steps with the reference's contents; config.yml; topo.nc"""
    from scipy.io import netcdf_file
    from oracle.ref_problem import FileIONetCDF
    text = SOMMERFELD.replace("tol: 1e-8", f"tol: {tol}").replace("max_it: 10_000", f"max_it: {max_it}") \
        .replace("write_freq: 1000", f"write_freq: {wf}").replace("silent: True", "silent: False")
    outs = []
    for k, make in enumerate((lambda t: Host.from_string(t), lambda t: _ref(ref, t, silent=False, output=str(tmp_path / 'r' / 'run')))):
        t = text.replace("output: data/journal", f"output: {tmp_path / 'a' / 'run'}\n    use_tstamp: False")
        FileIONetCDF.opened.clear()
        with contextlib.redirect_stdout(io.StringIO()):
            p = make(t)
            p.run()
        outs.append(p)
    a, r = outs
    assert a.step == r.step and (a.step < max_it) == (tol > 1e-6)
    assert np.array_equal(a.q, r.q) and a.dt == r.dt and a.simtime == r.simtime
    for key in r.history:
        assert np.array_equal(np.array(a.history[key], dtype=float), np.array(r.history[key], dtype=float)), key
    topo_r, sol_r = FileIONetCDF.opened
    assert sol_r.closed
    with netcdf_file(os.path.join(a.outdir, 'sol.nc'), 'r', mmap=False) as f:
        sol = {k: np.array(v[:]) for k, v in f.variables.items()}
    assert sol['solution'].shape[0] == len(sol_r.frames) == len(r.history['step'])
    for i, fr in enumerate(sol_r.frames):
        assert np.array_equal(sol['solution'][i, :, 0], fr['solution']), i
        assert np.array_equal(sol['pressure'][i], fr['pressure']), i
        assert np.array_equal(sol['wall_stress_xz'][i, :, 0], fr['wall_stress_xz']), i
        assert np.array_equal(sol['wall_stress_yz'][i, :, 0], fr['wall_stress_yz']), i
    with netcdf_file(os.path.join(a.outdir, 'topo.nc'), 'r', mmap=False) as f:
        assert np.array_equal(np.array(f.variables['topography'][:])[0, :, 0], topo_r.frames[0]['topography'])
    rows = {}
    for who, d in (('a', a.outdir), ('r', r.outdir)):
        with open(os.path.join(d, 'history.csv')) as f:
            rows[who] = list(csv.reader(f))
    assert rows['a'][0] == rows['r'][0] == ['step', 'time', 'ekin', 'residual', 'vsound']
    assert np.array_equal(np.array(rows['a'][1:], dtype=float), np.array(rows['r'][1:], dtype=float))
    import yaml
    ca, cr = (yaml.safe_load(open(os.path.join(d, 'config.yml'))) for d in (a.outdir, r.outdir))
    assert set(cr) - {'version'} == set(ca) - {'version'}
    for key in ('grid', 'numerics', 'geo', 'prop'):
        assert ca[key] == cr[key], key


def test_wave_decay_script_runs_unchanged(ref, Host):
    """the body of tests/test_wave_decay.py:85-108 (writes into q, assigns kinetic_energy_old, reads simtime) on both"""
    out = []
    for make in (Host.from_string, lambda t: _ref(ref, t)):
        with contextlib.redirect_stdout(io.StringIO()):
            p = make(DECAY)
            p._pre_run()
        kn = 2. * 2. * np.pi / p.grid['Lx']
        x = p.topo.x[1:-1, 1]
        p.q[2, 1:-1, :] = np.sin(kn * x)[:, None]
        p.kinetic_energy_old = p.kinetic_energy
        for _ in range(20):
            p.update()
        out.append((p.q.copy(), p.simtime, p.residual))
    assert np.array_equal(out[0][0], out[1][0]) and out[0][1:] == out[1][1:]


def test_mass_conservation_script(ref, Host):
    a, r = Host.from_string(JOURNAL2D.format(n=20)), _ref(ref, JOURNAL2D.format(n=20))
    for p in (a, r):
        with contextlib.redirect_stdout(io.StringIO()):
            p._pre_run()
    m0 = a.mass.copy()                                     # tests/test_mass_conservation.py:71 calls .copy() on it
    for _ in range(10):
        a.update()
        r.update()
    assert a.mass == r.mass and np.isclose(a.mass, m0)


# ---------------------------------------------------------------------------------------------------------------------
# GP-coupled step: SurrogateDriver / SurrogateModel / Database of gapflow_b200/surrogate.py (host logic) on doubles of the
# device pieces, against the executed reference with the stated tinygp stand-in
# ---------------------------------------------------------------------------------------------------------------------
@pytest.fixture
def HostGP(monkeypatch):
    import gapflow_b200.problem as P
    import gapflow_b200.surrogate as S
    from tests import host_double as HD
    monkeypatch.setattr(P, 'FVSolver', HD.OracleSurrogateSolver)
    monkeypatch.setattr(S, 'GPModel', HD.OracleGPModel)
    monkeypatch.setattr(S, '_features', HD.features)
    monkeypatch.setattr(S, '_capi', HD.CapiShim())
    return P.Problem


FIELD = {'zz': 'pressure', 'xz': 'wall_stress_xz', 'yz': 'wall_stress_yz'}


def _gp_pair(ref, HostGP, text, nx, ny):
    from gapflow_b200 import config
    from gapflow_b200.surrogate import Database
    from oracle import surrogate as OS
    from oracle import tinygp_stand_in as T
    from tests import surrogate_cases as SC
    extra = SC.slip_length(nx, ny)
    d = config.read_yaml_string(text)
    with io.StringIO(text) as f, contextlib.redirect_stdout(io.StringIO()):
        rd = ref.read_yaml_input(f)
    rd['options']['silent'] = True
    rd['db']['dtool_path'] = None
    T.HOOKS['draws'] = SC.DRAWS
    T.HOOKS['optimizer'] = lambda p0, X=None, **k: dict(zip(('log_amp', 'log_scale'),
                                                            SC.fixed_optimizer(type('M', (), {'Xtrain': X}))))
    database = Database(OS.MockMD(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS), d['db'])
    a = HostGP(d['options'], d['grid'], d['numerics'], d['properties'], d['geometry'], d['gp'], database, extra_field=extra)
    for m in a.gp_models.values():
        m.optimizer = SC.fixed_optimizer
    with contextlib.redirect_stdout(io.StringIO()):
        rdb = ref.Database(ref.Mock(rd['properties'], rd['geometry'], rd['gp']), rd['db'])
        r = ref.Problem(rd['options'], rd['grid'], rd['numerics'], rd['properties'], rd['geometry'], gp=rd['gp'],
                        database=rdb, extra_field=extra)
        for p in (a, r):
            p.q[...] = SC.perturb(p.q)
            p._communicate_ghost_buffers()
            p._pre_run()
    return a, database, r, rdb


def _rel(x, y):
    return np.abs(np.asarray(x) - np.asarray(y)).max() / max(np.abs(np.asarray(y)).max(), 1e-300)


def _same_gp(a, adb, r, rdb, tag, tol):
    assert adb.size == rdb.size, tag
    assert _rel(adb._Xtrain, rdb._Xtrain) <= tol and _rel(adb._Ytrain, rdb._Ytrain) <= max(tol, 1e-11), tag
    assert _rel(a.q, r.q) <= tol and abs(a.dt - r.dt) <= tol * r.dt and a.step == r.step, tag
    assert abs(a.residual - r.residual) <= 1e-9 * max(abs(r.residual), 1e-300), tag
    assert abs(a.pressure.v_sound - float(r.pressure.v_sound)) <= 1e-9 * float(r.pressure.v_sound), tag
    assert _rel(a.pressure.pressure, r.pressure.pressure) <= tol, tag
    assert _rel(a.wall_stress_xz.full, r.wall_stress_xz.full) <= tol and _rel(a.wall_stress_yz.full, r.wall_stress_yz.full) <= tol, tag
    assert set(a.gp_models) == {k for k, f in FIELD.items() if getattr(r, f).is_gp_model}, tag
    for k, m in a.gp_models.items():
        rm = getattr(r, FIELD[k])
        unit = m.kernel_variance * m.Yscale**2
        assert m.last_fit_train_size == rm.last_fit_train_size and m._pause == rm._pause and m._step == rm._step, (tag, k)
        assert np.abs(m.variance.numpy() - np.asarray(rm._predictive_var)).max() <= 1e-9 * unit, (tag, k)
        name = FIELD[k] + '_var'
        assert np.abs(a._variance_fields()[name] - rm.variance).max() <= 1e-9 * unit, (tag, k)      # what sol.nc gets
        assert abs(m.maximum_variance - rm.maximum_variance) <= 1e-9 * unit and m.trusted == bool(rm.trusted), (tag, k)
        assert abs(m.variance_tol - rm.variance_tol) <= max(tol, 1e-11) * rm.variance_tol, (tag, k)
        for key, hist in m.history.items():
            assert len(hist) == len(rm.history[key]) and (not hist or _rel(np.array(hist, dtype=float), rm.history[key]) <= 1e-9), (tag, k, key)


@pytest.mark.parametrize('name, kw, steps', [
    ('slip1d', dict(nx=100, atol=0.6, al_steps=3), 8),
    ('slip2d', dict(nx=24, ny=20, atol=0.8, al_steps=2), 4),
    ('slip1d', dict(nx=60, atol=0.05, al_steps=2), 6),
])
def test_gp_coupled_host_logic_like_the_reference(ref, HostGP, name, kw, steps):
    from tests import surrogate_cases as SC
    d, text = SC.inputs(name, **kw)
    a, adb, r, rdb = _gp_pair(ref, HostGP, text, d['grid']['Nx'], d['grid']['Ny'])
    _same_gp(a, adb, r, rdb, 'pre_run', 1e-12)
    grew = False
    for k in range(steps):
        n = adb.size
        with contextlib.redirect_stdout(io.StringIO()):
            a.update()
            r.update()
        _same_gp(a, adb, r, rdb, k, 1e-12)
        grew = grew or adb.size > n
    assert grew


@pytest.mark.parametrize('block', range(4))
def test_gp_coupled_host_logic_random_configurations(ref, HostGP, block):
    from tests.test_surrogate_oracle_vs_reference import _random_text
    for seed in range(block * 5, block * 5 + 5):
        text, nx, ny = _random_text(seed)
        a, adb, r, rdb = _gp_pair(ref, HostGP, text, nx, ny)
        _same_gp(a, adb, r, rdb, (seed, 'pre_run'), 1e-9)
        for k in range(5):
            with contextlib.redirect_stdout(io.StringIO()):
                a.update()
                r.update()
            _same_gp(a, adb, r, rdb, (seed, k), 1e-9)


def test_gp_run_outputs_like_the_reference(ref, HostGP, tmp_path):
    """run() with GP closures and outputs: frames of sol.nc (variance fields included, zeros in frame 0), history.csv,
    gp_zz.csv / gp_xz.csv (hyper-parameter history rows at the reference's steps), Xtrain.npy / Ytrain.npy"""
    from scipy.io import netcdf_file
    from gapflow_b200 import config
    from gapflow_b200.surrogate import Database
    from oracle import surrogate as OS
    from oracle import tinygp_stand_in as T
    from oracle.ref_problem import FileIONetCDF
    from tests import surrogate_cases as SC
    T.HOOKS['draws'] = SC.DRAWS
    T.HOOKS['optimizer'] = lambda p0, X=None, **k: dict(zip(('log_amp', 'log_scale'),
                                                            SC.fixed_optimizer(type('M', (), {'Xtrain': X}))))
    extra = SC.slip_length(64, 1)
    outs = {}
    for who in ('a', 'r'):
        _, text = SC.inputs('slip1d', nx=64, max_it=12, wf=5, out=str(tmp_path / who / 'run'), silent=False, atol=0.8, al_steps=2)
        FileIONetCDF.opened.clear()
        with contextlib.redirect_stdout(io.StringIO()):
            if who == 'a':
                d = config.read_yaml_string(text)
                db = Database(OS.MockMD(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS), d['db'])
                p = HostGP(d['options'], d['grid'], d['numerics'], d['properties'], d['geometry'], d['gp'], db, extra_field=extra)
                for m in p.gp_models.values():
                    m.optimizer = SC.fixed_optimizer
            else:
                with io.StringIO(text) as f:
                    rd = ref.read_yaml_input(f)
                rd['db']['dtool_path'] = None
                db = ref.Database(ref.Mock(rd['properties'], rd['geometry'], rd['gp']), rd['db'])
                p = ref.Problem(rd['options'], rd['grid'], rd['numerics'], rd['properties'], rd['geometry'], gp=rd['gp'],
                                database=db, extra_field=extra)
            p.run()
        outs[who] = (p, db)
    (a, adb), (r, rdb) = outs['a'], outs['r']
    assert a.step == r.step == 12 and adb.size == rdb.size > 10
    sol_r = FileIONetCDF.opened[-1]
    with netcdf_file(os.path.join(a.outdir, 'sol.nc'), 'r', mmap=False) as f:
        sol = {k: np.array(v[:]) for k, v in f.variables.items()}
    assert sol['solution'].shape[0] == len(sol_r.frames) == 4               # steps 0, 5, 10, 12
    for i, fr in enumerate(sol_r.frames):
        assert _rel(sol['solution'][i, :, 0], fr['solution']) <= 1e-12, i
        for name in ('pressure', 'pressure_var', 'wall_stress_xz_var'):
            assert np.abs(sol[name][i] - fr[name]).max() <= 1e-9 * max(np.abs(fr[name]).max(), 1e-300), (i, name)
        assert _rel(sol['wall_stress_xz'][i, :, 0], fr['wall_stress_xz']) <= 1e-12, i
    assert not sol['pressure_var'][0].any() and sol['pressure_var'][1].any()
    for name in ('history.csv', 'gp_zz.csv', 'gp_xz.csv'):
        rows = []
        for p in (a, r):
            with open(os.path.join(p.outdir, name)) as f:
                rows.append(list(csv.reader(f)))
        assert rows[0][0] == rows[1][0], name                                # the same columns in the same order
        assert len(rows[0]) == len(rows[1]) > 1, name
        A, R = np.array(rows[0][1:], dtype=float), np.array(rows[1][1:], dtype=float)
        assert np.allclose(A, R, rtol=1e-9, atol=0), name
    for name in ('Xtrain.npy', 'Ytrain.npy', 'Ytrain_err.npy'):
        assert _rel(np.load(os.path.join(a.outdir, name)), np.load(os.path.join(r.outdir, name))) <= 1e-11, name
