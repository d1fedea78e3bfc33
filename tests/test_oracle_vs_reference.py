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


def test_piezoviscosity_laws_match_reference():
    """oracle/closures.py::piezoviscosity vs GaPFlow/models/viscosity.py:34-66 (all four laws + none)."""
    from oracle import closures as C
    R = ref_loader.modules()['viscosity']
    rng = np.random.default_rng(11)
    p = rng.uniform(1e5, 5e8, (13, 7))
    rho = rng.uniform(1.0, 849.0, (13, 7))
    cases = [(p, dict(name='Barus', aB=2e-8)), (p, dict(name='Roelands', mu_inf=1e-3, p_ref=1.96e8, z=0.68)),
             (rho, dict(name='Dukler', eta_v=3.9e-5, rho_l=850., rho_v=0.019)),
             (rho, dict(name='McAdams', eta_v=3.9e-5, rho_l=850., rho_v=0.019)), (p, dict(name='none'))]
    for x, pz in cases:
        assert np.array_equal(C.piezoviscosity(x, 0.039, pz), R.piezoviscosity(x, 0.039, dict(pz))), pz['name']
    g = rng.uniform(-1e9, 1e9, (13, 7))
    h = rng.uniform(1e-6, 1e-4, (13, 7))
    assert np.array_equal(C.shear_rate_avg(g, 0.5 * g, h, 0.1, 0., 0.039), R.shear_rate_avg(g, 0.5 * g, h, 0.1, 0., 0.039))
