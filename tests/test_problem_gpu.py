"""GPU: the reference's own integration tests, run through the drop-in ``Problem``
(gapflow_b200/problem.py).  YAML texts are the reference's (tests/test_sommerfeld.py:32-68,
tests/test_mass_conservation.py:30-64, tests/test_flip_axes.py:30-65)."""
import io
import os
from copy import deepcopy

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

from gapflow_b200 import config  # noqa: E402

_PROP = """
properties:
    shear: 0.0794
    bulk: 0.
    EOS: DH
    P0: 101325.
    rho0: 877.7007
    T0: 323.15
    C1: {C1}
    C2: 1.23
"""
_PER = "    xE: ['P', 'P', 'P']\n    xW: ['P', 'P', 'P']\n    yS: ['P', 'P', 'P']\n    yN: ['P', 'P', 'P']\n"
SOMMERFELD = ("options:\n    output: data/journal\n    write_freq: 1000\n    silent: True\n"
              "grid:\n    dx: 1.e-5\n    dy: 1.\n    Nx: 100\n    Ny: 1\n" + _PER +
              "geometry:\n    type: journal\n    CR: 1.e-2\n    eps: 0.7\n    U: 0.1\n    V: 0.\n"
              "numerics:\n    CFL: 0.5\n    adaptive: 1\n    tol: 1e-8\n    dt: 1e-10\n    max_it: 10_000\n"
              + _PROP.format(C1='3.5e12'))
JOURNAL2D = ("options:\n    output: data/journal\n    write_freq: 1000\n    silent: True\n"
             "grid:\n    dx: 1.e-5\n    dy: 1.e-5\n    Nx: {n}\n    Ny: {n}\n" + _PER +
             "geometry:\n    type: journal\n    CR: 1.e-2\n    eps: 0.7\n    U: 0.1\n    V: 0.\n"
             "numerics:\n    CFL: 0.5\n    adaptive: 1\n    tol: 1e-8\n    dt: 1e-10\n    max_it: 5000\n"
             + _PROP.format(C1='3.5e10'))


def sommerfeld_solution(x, Lx, mu, U, clearance_ratio, eps, P0):
    Rb = Lx / (2. * np.pi)
    c = clearance_ratio * Rb
    omega = U / Rb
    prefac = 6. * mu * omega * (Rb / c)**2 * eps
    return P0 + prefac * np.sin(x / Rb) * (2. + eps * np.cos(x / Rb)) / ((2. + eps**2) * (1. + eps * np.cos(x / Rb))**2)


@pytest.mark.parametrize('eps', [0.5, 0.7, 0.9])
def test_pressure_profile_sommerfeld(eps):
    """tests/test_sommerfeld.py:115-141."""
    from gapflow_b200.problem import Problem
    with io.StringIO(SOMMERFELD) as f:
        d = config.read_yaml_input(f)
    d['geometry']['eps'] = eps
    problem = Problem._from_dict(d)
    problem.run()
    p_num = problem.pressure.pressure[1:-1, 1]
    Lx = problem.grid['Lx']
    x_ana = np.linspace(0., Lx, 101)
    x_num = (x_ana[1:] + x_ana[:-1]) / 2.
    dp = p_num[1] - p_num[0]
    p_ana = sommerfeld_solution(x_num, Lx, problem.pressure.prop['shear'], problem.pressure.geo['U'],
                                problem.pressure.geo['CR'], eps, p_num[0] - dp / 2)
    assert np.linalg.norm(p_ana - p_num) / np.linalg.norm(p_ana) < 0.02
    assert problem.step > 100


def test_mass_conservation():
    """tests/test_mass_conservation.py:67-77."""
    from gapflow_b200.problem import Problem
    problem = Problem.from_string(JOURNAL2D.format(n=50))
    problem._pre_run()
    mass_before = problem.mass.copy()
    for _ in range(50):
        problem.update()
    assert np.isclose(problem.mass, mass_before)
    assert problem.step == 50


def test_flip_axes():
    """tests/test_flip_axes.py:68-96."""
    from gapflow_b200.problem import Problem
    with io.StringIO(JOURNAL2D.format(n=100)) as f:
        input_x = config.read_yaml_input(f)
    input_y = deepcopy(input_x)
    input_y['geometry']['U'] = 0.
    input_y['geometry']['V'] = input_x['geometry']['U']
    input_y['geometry']['flip'] = True
    px, py = Problem._from_dict(input_x), Problem._from_dict(input_y)
    px._pre_run()
    py._pre_run()
    for _ in range(5):
        px.update()
        py.update()
        np.testing.assert_almost_equal(px.q[0, 1:-1, 1:-1], py.q[0, 1:-1, 1:-1].T)
        np.testing.assert_almost_equal(px.q[1, 1:-1, 1:-1], py.q[2, 1:-1, 1:-1].T)
        np.testing.assert_almost_equal(px.q[2, 1:-1, 1:-1], py.q[1, 1:-1, 1:-1].T)


def test_host_edits_of_q_are_honoured_and_match_oracle():
    """User code writes problem.q[...] between steps (tests/test_wave_decay.py:101)."""
    from gapflow_b200.problem import Problem
    from oracle.fv import FVOracle
    d = config.read_yaml_string(JOURNAL2D.format(n=24))
    problem = Problem._from_dict(d)
    o = FVOracle(d['grid'], d['geometry'], d['properties'], d['numerics'])
    rng = np.random.default_rng(3)
    bump = 1 + 1e-6 * rng.standard_normal(o.q[0].shape)
    problem.q[0] *= bump
    o.q[0] *= bump
    problem._pre_run()
    o.dt = d['numerics']['CFL'] * o.dt_crit
    assert problem.dt == o.dt
    for k in range(3):
        problem.update()
        o.update()
        assert np.array_equal(problem.q, o.q)
    problem.q[1] += 1e-4           # edit in place mid-run, no _pre_run
    o.q[1] += 1e-4
    problem.update()
    o.update()
    assert np.array_equal(problem.q, o.q) and problem.dt == o.dt
    # stale closure fields, as the reference stores them: those of the corrector pass's input
    assert np.array_equal(problem.pressure.pressure, o.pressure)
    assert np.array_equal(problem.wall_stress_xz.full, o.wall_xz)
    assert np.array_equal(problem.bulk_stress.stress, o.bulk)


def test_run_writes_the_reference_files(tmp_path, capsys):
    from scipy.io import netcdf_file
    from gapflow_b200.problem import Problem
    d = config.read_yaml_string(SOMMERFELD.replace('C1: 3.5e12', 'C1: 3.5e10'))
    d['options'].update(output=str(tmp_path / 'run'), silent=False, use_tstamp=False, write_freq=100)
    d['numerics']['max_it'] = 250
    problem = Problem._from_dict(d)
    problem.run()
    out = tmp_path / 'run'
    assert sorted(os.listdir(out)) == ['config.yml', 'history.csv', 'sol.nc', 'topo.nc']
    rows = (out / 'history.csv').read_text().strip().split('\n')
    assert rows[0] == 'step,time,ekin,residual,vsound'
    assert [int(r.split(',')[0]) for r in rows[1:]] == [0, 100, 200, 250]
    with netcdf_file(str(out / 'sol.nc'), 'r', mmap=False) as f:
        # muGrid's layout as the reference's readers index it (viz/plotting.py:331-349): (frame, comp, sub-point, x, y)
        assert f.variables['solution'].shape == (4, 3, 1, 102, 3)
        assert f.variables['wall_stress_xz'].shape == (4, 12, 1, 102, 3)
        assert f.variables['pressure'].shape == (4, 102, 3)
        q_nc = np.asarray(f.variables['solution'][:])[-1]
        assert np.array_equal(q_nc[:, 0], problem.q)
        assert 'pressure_var' not in f.variables
    with netcdf_file(str(out / 'topo.nc'), 'r', mmap=False) as f:
        topo = np.asarray(f.variables['topography'][:])
        assert np.array_equal(topo[0, 0, 0, 1:-1, 1], problem.topo.h[1:-1, 1])      # plotting.py:235
    assert "steps/s" in capsys.readouterr().out


# ---------------------------------------------------------------------------------------------
# tests/test_wave_decay.py:31-146: decay of a shear wave and of a sound wave in a flat channel
# (cubic EOS, fixed dt, host edits of problem.q before the first update)
# ---------------------------------------------------------------------------------------------
DECAY = ("options:\n    output: data/decay\n    write_freq: 100\n    use_tstamp: False\n    silent: True\n"
         "grid:\n    Lx: 3.2e-7\n    Ly: 1\n    Nx: 256\n    Ny: 1\n" + _PER +
         "geometry:\n    type: inclined\n    hmin: 5e-9\n    hmax: 5e-9\n    U: 0.\n    V: 0.\n"
         "numerics:\n    adaptive: 0\n    CFL: 0.5\n    dt: 1e-13\n    max_it: 5_000\n"
         "properties:\n    EOS: cubic\n    shear: 3.92293e-05\n    bulk: 0.\n    rho0: 762.8617\n"
         "    a: 1.33030e-1\n    b: -1.41778e2\n    c: 8.35134e4\n    d: -2.86532e6\n")


def _decay_problem():
    from gapflow_b200.problem import Problem
    with io.StringIO(DECAY) as f:
        d = config.read_yaml_input(f)
    problem = Problem._from_dict(d)
    problem._pre_run()
    return problem


@pytest.mark.parametrize('n', [1, 2, 4, 8])
def test_shear_wave_decay(n):
    """tests/test_wave_decay.py:85-108."""
    problem = _decay_problem()
    h = problem.geo['hmin']
    kin_visc = problem.prop['shear'] / problem.prop['rho0']
    kn = n * 2. * np.pi / problem.grid['Lx']
    tau = h**2 / (6 * kin_visc)
    x = problem.topo.x[1:-1, 1]
    problem.q[2, 1:-1, :] = np.sin(kn * x)[:, None]
    problem.kinetic_energy_old = problem.kinetic_energy
    for i in range(200):
        problem.update()
        jy_ana = np.sin(kn * x) * np.exp(-2 * problem.simtime / tau)
        np.testing.assert_almost_equal(problem.q[2, 1:-1, 1], jy_ana, decimal=4)


@pytest.mark.parametrize('n', [1, 2, 3, 4])
def test_sound_wave_decay(n):
    """tests/test_wave_decay.py:111-146."""
    problem = _decay_problem()
    h = problem.geo['hmin']
    kin_visc = problem.prop['shear'] / problem.prop['rho0']
    kn = n * 2. * np.pi / problem.grid['Lx']
    tau = h**2 / (6 * kin_visc)
    cT = problem.pressure.v_sound
    x = problem.topo.x[1:-1, 1]
    problem.q[1, 1:-1, :] = np.sin(kn * x)[:, None]
    problem.kinetic_energy_old = problem.kinetic_energy
    k_crit = 6. * kin_visc / (h**2 * cT)
    for i in range(400):
        problem.update()
        t = problem.simtime
        if kn > k_crit:
            sT = np.sqrt(cT**2 - (1 / tau / kn)**2)
            env = np.exp(-t / tau) * (np.cos(sT * kn * t) - np.sin(sT * kn * t) / (tau * sT * kn))
        else:
            isT = np.sqrt((1 / tau / kn)**2 - cT**2)
            env = np.exp(-t / tau) * (np.cosh(isT * kn * t) - np.sinh(isT * kn * t) / (tau * isT * kn))
        np.testing.assert_almost_equal(problem.q[1, 1:-1, 1], np.sin(kn * x) * env, decimal=3)


# ---------------------------------------------------------------------------------------------
# Problem.run() on x-slabs (gapflow_b200.dist.Parallel): same files as the single-domain run
# ---------------------------------------------------------------------------------------------
def _run_problem_threads(make, world, collect=None):
    """one Problem per rank, ranks = threads of this process on one GPU (ThreadColl stands in for torch.distributed;
    the launches of all ranks go down one stream in lock step, so no kernel ever waits for a peer).  ``collect(p)``
    runs on every rank's thread after run(): q, mass, ... of a slab-decomposed Problem are GLOBAL quantities whose
    access is a collective of all ranks."""
    import threading
    from gapflow_b200.dist import Parallel
    from gapflow_b200.surrogate import ThreadColl
    tc = ThreadColl(world)
    out, errs = [None] * world, []

    def work(r):
        try:
            torch.cuda.set_device(0)
            p = make(Parallel.threads(tc, r))
            p.run()
            out[r] = collect(p) if collect else p
        except BaseException as e:
            import traceback
            errs.append((r, traceback.format_exc()))
            tc._barrier.abort()

    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join(600)
    assert not errs, errs
    return out


def _read_nc(path):
    from scipy.io import netcdf_file
    with netcdf_file(path, 'r', mmap=False) as f:
        return {k: np.array(v[:]) for k, v in f.variables.items()}


@pytest.mark.parametrize("world", [2, 3])
def test_problem_run_on_slabs_writes_the_same_files(tmp_path, world, capsys):
    from gapflow_b200.problem import Problem
    text = SOMMERFELD.replace('C1: 3.5e12', 'C1: 3.5e10')

    def make(par, tag):
        d = config.read_yaml_string(text)
        d['options'].update(output=str(tmp_path / tag), silent=False, use_tstamp=False, write_freq=50)
        d['numerics']['max_it'] = 120
        return Problem(d['options'], d['grid'], d['numerics'], d['properties'], d['geometry'], parallel=par)
    one = make(None, 'one')
    one.run()
    ranks = _run_problem_threads(lambda par: make(par, f'slabs{world}'), world,
                                 collect=lambda p: (p.step, p.dt, p.q.copy(), p.mass, p.kinetic_energy))
    assert all(st == one.step == 120 for st, *_ in ranks)
    assert all(dt == one.dt and np.array_equal(q, one.q) for _, dt, q, _, _ in ranks)      # q, dt: global and bit-identical
    assert all(abs(m - one.mass) <= 1e-12 * one.mass for *_, m, _ in ranks)
    assert all(abs(e - one.kinetic_energy) <= 1e-12 * one.kinetic_energy for *_, e in ranks)
    a, b = _read_nc(str(tmp_path / 'one' / 'sol.nc')), _read_nc(str(tmp_path / f'slabs{world}' / 'sol.nc'))
    assert sorted(a) == sorted(b)
    for k in a:
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(_read_nc(str(tmp_path / 'one' / 'topo.nc'))['topography'],
                          _read_nc(str(tmp_path / f'slabs{world}' / 'topo.nc'))['topography'])
    ha = np.loadtxt(str(tmp_path / 'one' / 'history.csv'), delimiter=',', skiprows=1)
    hb = np.loadtxt(str(tmp_path / f'slabs{world}' / 'history.csv'), delimiter=',', skiprows=1)
    assert ha.shape == hb.shape and np.array_equal(ha[:, :2], hb[:, :2]) and np.array_equal(ha[:, 4], hb[:, 4])
    assert np.allclose(ha[:, 2:4], hb[:, 2:4], rtol=1e-9, atol=0)      # the kinetic-energy sum is folded rank by rank
    assert sorted(os.listdir(tmp_path / f'slabs{world}')) == ['config.yml', 'history.csv', 'sol.nc', 'topo.nc']
    capsys.readouterr()
