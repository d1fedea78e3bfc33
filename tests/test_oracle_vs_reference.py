"""CPU, build container only: oracle vs the reference's own NumPy functions imported from
/root/reference (skipped where the reference is absent, e.g. on the GPU box)."""
import numpy as np
import pytest

from oracle import closures as C
from oracle import ref_loader
from oracle.make_golden import EOS_PROPS, EOS_DENS, RefDrivenOracle
from tests.cases import STEP_CASES, build_case

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")


@pytest.fixture(scope="module")
def R():
    return ref_loader.modules()


def test_random_closures(R):
    rng = np.random.default_rng(99)
    for trial in range(20):
        shape = (11, 13)
        q = np.stack([800 + 100 * rng.random(shape), 50 * rng.standard_normal(shape), 50 * rng.standard_normal(shape)])
        h = np.stack([1e-5 * (0.2 + rng.random(shape)), 1e-2 * rng.standard_normal(shape), 1e-2 * rng.standard_normal(shape)])
        Ls = 1e-5 * rng.random((1, *shape)) * (trial % 2)
        U, V, eta, zeta = rng.standard_normal(), rng.standard_normal() * (trial % 3 > 0), rng.random(), rng.random() * (trial % 4 > 1)
        for fn in ("stress_bottom", "stress_top", "stress_avg"):
            a = getattr(R['viscous'], fn)(q, h, U, V, eta, zeta, Ls)
            b = getattr(C, fn)(q, h, U, V, eta, zeta, Ls[0])
            assert np.array_equal(a, b), fn


def test_eos(R):
    rng = np.random.default_rng(5)
    for p in EOS_PROPS:
        mu, w = EOS_DENS[p['EOS']]
        d = mu * (1 + w * (rng.random(1000) - 0.5) * 2)
        assert np.array_equal(R['pressure'].eos_pressure(d.copy(), p), C.eos_pressure(d.copy(), p))
        with np.errstate(invalid='ignore'):
            assert np.array_equal(R['sound'].eos_sound_velocity(d.copy(), p), C.eos_sound_velocity(d.copy(), p),
                                  equal_nan=True)


@pytest.mark.parametrize("name", list(STEP_CASES))
def test_ten_steps(R, name):
    RefDrivenOracle.R = R
    a = build_case(name, cls=RefDrivenOracle)
    b = build_case(name)
    for _ in range(10):
        a.update()
        b.update()
    assert np.array_equal(a.q, b.q)
    assert a.dt == b.dt and a.residual == b.residual
