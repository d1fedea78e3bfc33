"""CPU: the NumPy oracle reproduces, bit for bit, the golden vectors that
oracle/make_golden.py generated from the reference's own functions."""
import os

import numpy as np
import pytest

from oracle import closures as C
from oracle.make_golden import EOS_PROPS
from tests.cases import STEP_CASES, build_case


@pytest.fixture(scope="module")
def gc(golden_dir):
    return np.load(os.path.join(golden_dir, "fv_closures.npz"))


@pytest.fixture(scope="module")
def gs(golden_dir):
    return np.load(os.path.join(golden_dir, "fv_steps.npz"))


@pytest.mark.parametrize("fn", ["stress_bottom", "stress_top", "stress_avg"])
def test_stress_bit_exact(gc, fn):
    q, h = gc['q'], gc['h']
    for k, (U, V, eta, zeta) in enumerate(gc['visc_params']):
        for ls_name, ls in (('ls0', np.zeros_like(gc['Ls'])), ('ls', gc['Ls'])):
            got = getattr(C, fn)(q, h, U, V, eta, zeta, ls)
            assert np.array_equal(got, gc[f'{fn}_{k}_{ls_name}'])


@pytest.mark.parametrize("prop", EOS_PROPS, ids=[p['EOS'] for p in EOS_PROPS])
def test_eos_bit_exact(gc, prop):
    d = gc[f"dens_{prop['EOS']}"]
    p, wp = C.eos_pressure(d.copy(), prop), gc[f"p_{prop['EOS']}"]
    with np.errstate(invalid='ignore'):
        c, wc = C.eos_sound_velocity(d.copy(), prop), gc[f"c_{prop['EOS']}"]
    if prop['EOS'] in ('DH', 'vdW'):      # + - * / sqrt only: machine independent
        assert np.array_equal(p, wp)
        assert np.array_equal(c, wc, equal_nan=True)
    else:                                  # np.power/exp/log are CPU-dispatched SIMD routines (~1 ulp)
        assert np.allclose(p, wp, rtol=1e-14, atol=1e-14 * np.abs(wp).max())
        assert np.allclose(c, wc, rtol=1e-14, equal_nan=True)


@pytest.mark.parametrize("name", list(STEP_CASES))
def test_steps_bit_exact(gs, name):
    if 'EOS: BWR' in STEP_CASES[name]['yaml']:
        pytest.skip("BWR uses np.power/np.exp (CPU-dispatched SIMD): covered with a tolerance below")
    o = build_case(name)
    assert np.array_equal(o.q, gs[f'{name}_q0'])
    assert o.dt == gs[f'{name}_dt0']
    o.update()
    assert np.array_equal(o.q, gs[f'{name}_q1'])
    s = np.array([o.dt, o.residual, o.kinetic_energy, o.v_max, o.v_sound, o.mass])
    assert np.array_equal(s, gs[f'{name}_s1'])
    o.update()
    o.update()
    assert np.array_equal(o.q, gs[f'{name}_q3'])
    s = np.array([o.dt, o.residual, o.kinetic_energy, o.v_max, o.v_sound, o.mass])
    assert np.array_equal(s, gs[f'{name}_s3'])


def test_bwr_steps_within_tolerance(gs):
    name = 'bwr2d_slip'
    o = build_case(name)
    for n, tag in ((1, '1'), (2, '3')):
        for _ in range(n):
            o.update()
        want = gs[f'{name}_q{tag}']
        assert np.abs(o.q - want).max() <= 1e-13 * np.abs(want).max()
