"""The general closure signatures stress_bottom / stress_top / stress_avg(..., dqx, dqy, slip) (viscous.py:37, 281, 612):
slip at both walls and in-plane gradients.  Closed forms derived from the physics (tools/derive_slip_closures.py),
checked here against golden vectors generated from the reference's own functions (oracle/make_golden.py) --
on the CPU through the library's host instantiation of the same routine, on the GPU through gpf_stress_ex -- and,
where the reference tree is present, by the reference's analytic tests (tests/test_analytic.py:31-123)."""
import ctypes as C
import os

import numpy as np
import pytest

from gapflow_b200 import _capi
from gapflow_b200.fv import _pure_cfg

TOL = 1e-12       # of the largest component: the reference's expanded polynomials cancel where ours do not
FNS = {'stress_bottom': (0, 6), 'stress_top': (1, 6), 'stress_avg': (2, 3)}
VARIANTS = [('both', 'dq0'), ('both', 'dq'), ('top', 'dq')]


@pytest.fixture(scope="module")
def gc(golden_dir):
    return np.load(os.path.join(golden_dir, "fv_closures.npz"))


def _host(fn, q, h, U, V, eta, zeta, Ls, dqx, dqy, slip):
    lib = _capi.load()
    which, nc = FNS[fn]
    cfg = _pure_cfg(dict(EOS='DH', rho0=1., P0=0., C1=1., C2=2., shear=eta, bulk=zeta), dict(U=U, V=V))
    n = q[0].size
    arr = lambda a: None if a is None else np.ascontiguousarray(np.broadcast_to(a, (a.shape[0] if a.ndim == q.ndim else 1, *q.shape[1:])), dtype=np.float64)
    qa, ha = arr(q), arr(h[:3])
    la = None if Ls is None else np.ascontiguousarray(np.broadcast_to(Ls, q.shape[1:]), dtype=np.float64)
    gx, gy = arr(dqx), arr(dqy)
    out = np.empty((nc, *q.shape[1:]))
    ptr = lambda a: None if a is None else C.c_void_p(a.ctypes.data)
    _capi.check(lib.gpf_selftest_stress_ex_host(C.byref(cfg), which, _capi.SLIP[slip], ptr(qa), ptr(ha), ptr(la), ptr(gx), ptr(gy),
                                                ptr(out), n))
    return out


@pytest.mark.parametrize("fn", list(FNS))
def test_host_routine_vs_reference_golden(gc, fn):
    q, h, Ls = gc['q'], gc['h'], gc['Ls']
    for k, (U, V, eta, zeta) in enumerate(gc['visc_params']):
        for slip, gname in VARIANTS:
            gx, gy = (gc['dqx'], gc['dqy']) if gname == 'dq' else (None, None)
            want = gc[f'{fn}_{k}_{slip}_{gname}']
            got = _host(fn, q, h, float(U), float(V), float(eta), float(zeta), Ls, gx, gy, slip)
            assert np.abs(got - want).max() <= TOL * np.abs(want).max(), (fn, k, slip, gname)
        # the driver's call through the general routine: within tolerance of the bit-exact kernels' golden
        got = _host(fn, q, h, float(U), float(V), float(eta), float(zeta), Ls, None, None, 'top')
        want = gc[f'{fn}_{k}_ls']
        assert np.abs(got - want).max() <= TOL * np.abs(want).max(), (fn, k)


@pytest.mark.gpu
@pytest.mark.parametrize("fn", list(FNS))
def test_device_functions_vs_reference_golden(gc, fn):
    torch = pytest.importorskip("torch")
    from gapflow_b200 import fv
    dev = lambda x: torch.as_tensor(np.ascontiguousarray(x), device='cuda')
    q, h, Ls = dev(gc['q']), dev(gc['h']), dev(gc['Ls'])
    for k, (U, V, eta, zeta) in enumerate(gc['visc_params']):
        for slip, gname in VARIANTS:
            gx, gy = (dev(gc['dqx']), dev(gc['dqy'])) if gname == 'dq' else (None, None)
            got = getattr(fv, fn)(q, h, float(U), float(V), float(eta), float(zeta), Ls, dqx=gx, dqy=gy, slip=slip).cpu().numpy()
            want = gc[f'{fn}_{k}_{slip}_{gname}']
            assert np.abs(got - want).max() <= TOL * np.abs(want).max(), (fn, k, slip, gname)
            host = _host(fn, gc['q'], gc['h'], float(U), float(V), float(eta), float(zeta), gc['Ls'],
                         None if gx is None else gc['dqx'], None if gy is None else gc['dqy'], slip)
            assert np.abs(got - host).max() <= 1e-14 * np.abs(want).max()      # same routine, host and device
        # slip="top" without gradients keeps the bit-exact kernels
        got = getattr(fv, fn)(q, h, float(U), float(V), float(eta), float(zeta), Ls).cpu().numpy()
        assert np.array_equal(got, gc[f'{fn}_{k}_ls'])
    with pytest.raises(ValueError):
        fv.stress_avg(q, h, 0.1, 0., 1., 0., Ls, slip='bottom')


# ---- the reference's analytic tests (tests/test_analytic.py:58-123) against the library's functions ----------------
def _ref_profiles():
    from oracle import ref_loader
    if not ref_loader.available():
        pytest.skip("reference tree not present")
    return ref_loader.load("GaPFlow/models/profiles.py", "_ref_profiles")


@pytest.mark.parametrize('slip, Ls', [('both', 0.), ('both', 0.5), ('top', 0.), ('top', 0.5)])
def test_avg_and_wall_stress_match_the_integrated_profiles(slip, Ls):
    P = _ref_profiles()
    q_test = np.array([1.0, 0.75, 0.25])
    h_test = np.array([1.0, 0.01, 0.01])
    z = np.linspace(0., 1., 10_000)
    txx, tyy, tzz, tyz, txz, txy = P.get_stress_profiles(z, h_test, q_test, np.zeros(3), np.zeros(3), U=1., V=1., eta=1., zeta=1.,
                                                         Ls=Ls, mode=slip)
    q, h = q_test[:, None], h_test[:, None]
    ls = np.array([Ls])
    avg = _host('stress_avg', q, h, 1., 1., 1., 1., ls, None, None, slip)[:, 0]
    trap = getattr(np, 'trapezoid', getattr(np, 'trapz', None))
    assert np.isclose(trap(txx, z) / avg[0], 1.) and np.isclose(trap(tyy, z) / avg[1], 1.) and np.isclose(trap(txy, z) / avg[2], 1.)
    top = _host('stress_top', q, h, 1., 1., 1., 1., ls, None, None, slip)[:, 0]
    bot = _host('stress_bottom', q, h, 1., 1., 1., 1., ls, None, None, slip)[:, 0]
    for c, prof in enumerate((txx, tyy, tzz, tyz, txz, txy)):
        assert np.isclose(bot[c], prof[0]) and np.isclose(top[c], prof[-1]), (slip, Ls, c)
