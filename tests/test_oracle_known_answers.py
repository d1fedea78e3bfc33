"""CPU: the reference's own known-answer tests, run against the ORACLE (oracle/fv.py) -- the pin SURVEY 8(c) asks for
next to the golden vectors: what the reference asserts about its update loop, the restatement must satisfy too.
(tests/test_problem_gpu.py runs the same tests through the CUDA ``Problem``.)

  tests/test_sommerfeld.py:115-141        journal bearing pressure vs the Sommerfeld solution (< 2 % relative L2)
  tests/test_mass_conservation.py:67-77   periodic 2D journal, 50 steps
  tests/test_flip_axes.py:68-96           x/y symmetry to 7 decimals over 5 steps
  tests/test_wave_decay.py:85-146         decay of a shear wave (4 decimals, 200 steps) and a sound wave (3, 400 steps)
"""
from copy import deepcopy

import numpy as np
import pytest

from gapflow_b200 import config
from oracle.fv import FVOracle, midpoint_grid
from oracle import closures as C
from tests.test_problem_gpu import DECAY, JOURNAL2D, SOMMERFELD, sommerfeld_solution


def _oracle(d):
    return FVOracle(d['grid'], d['geometry'], d['properties'], d['numerics'])


@pytest.mark.parametrize('eps', [0.5, 0.7, 0.9])
def test_pressure_profile_sommerfeld(eps):
    d = config.read_yaml_string(SOMMERFELD)
    d['geometry']['eps'] = eps
    o = _oracle(d)
    o.run()
    assert o.step > 100 and not o._stop
    o.closures()
    p_num = o.pressure[1:-1, 1]
    Lx = d['grid']['Lx']
    x_ana = np.linspace(0., Lx, 101)
    x_num = (x_ana[1:] + x_ana[:-1]) / 2.
    dp = p_num[1] - p_num[0]
    p_ana = sommerfeld_solution(x_num, Lx, d['properties']['shear'], d['geometry']['U'], d['geometry']['CR'], eps,
                                p_num[0] - dp / 2)
    assert np.linalg.norm(p_ana - p_num) / np.linalg.norm(p_ana) < 0.02


def test_mass_conservation():
    o = _oracle(config.read_yaml_string(JOURNAL2D.format(n=50)))
    mass_before = o.mass
    for _ in range(50):
        o.update()
    assert o.step == 50
    assert np.isclose(o.mass, mass_before)


def test_flip_axes():
    dx = config.read_yaml_string(JOURNAL2D.format(n=100))
    dy = deepcopy(dx)
    dy['geometry']['U'] = 0.
    dy['geometry']['V'] = dx['geometry']['U']
    dy['geometry']['flip'] = True
    ox, oy = _oracle(dx), _oracle(dy)
    for _ in range(5):
        ox.update()
        oy.update()
        np.testing.assert_almost_equal(ox.q[0, 1:-1, 1:-1], oy.q[0, 1:-1, 1:-1].T)
        np.testing.assert_almost_equal(ox.q[1, 1:-1, 1:-1], oy.q[2, 1:-1, 1:-1].T)
        np.testing.assert_almost_equal(ox.q[2, 1:-1, 1:-1], oy.q[1, 1:-1, 1:-1].T)


def _decay():
    d = config.read_yaml_string(DECAY)
    o = _oracle(d)
    xx, _ = midpoint_grid(d['grid'])
    h = d['geometry']['hmin']
    kin_visc = d['properties']['shear'] / d['properties']['rho0']
    return d, o, xx[1:-1, 1], h**2 / (6 * kin_visc), kin_visc, h


@pytest.mark.parametrize('n', [1, 2, 4, 8])
def test_shear_wave_decay(n):
    d, o, x, tau, _, _ = _decay()
    kn = n * 2. * np.pi / d['grid']['Lx']
    o.q[2, 1:-1, :] = np.sin(kn * x)[:, None]
    o.kinetic_energy_old = o.kinetic_energy
    for _ in range(200):
        o.update()
        np.testing.assert_almost_equal(o.q[2, 1:-1, 1], np.sin(kn * x) * np.exp(-2 * o.simtime / tau), decimal=4)


@pytest.mark.parametrize('n', [1, 2, 3, 4])
def test_sound_wave_decay(n):
    d, o, x, tau, kin_visc, h = _decay()
    kn = n * 2. * np.pi / d['grid']['Lx']
    cT = C.eos_sound_velocity(o.q[0], d['properties']).max()
    o.q[1, 1:-1, :] = np.sin(kn * x)[:, None]
    o.kinetic_energy_old = o.kinetic_energy
    k_crit = 6. * kin_visc / (h**2 * cT)
    for _ in range(400):
        o.update()
        t = o.simtime
        if kn > k_crit:
            sT = np.sqrt(cT**2 - (1 / tau / kn)**2)
            env = np.exp(-t / tau) * (np.cos(sT * kn * t) - np.sin(sT * kn * t) / (tau * sT * kn))
        else:
            isT = np.sqrt((1 / tau / kn)**2 - cT**2)
            env = np.exp(-t / tau) * (np.cosh(isT * kn * t) - np.sinh(isT * kn * t) / (tau * isT * kn))
        np.testing.assert_almost_equal(o.q[1, 1:-1, 1], np.sin(kn * x) * env, decimal=3)
