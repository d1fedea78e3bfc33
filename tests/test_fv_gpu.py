"""GPU parity tests of the FV hot path: CUDA path (through the C ABI) vs the golden vectors
generated from the reference's own functions and vs the NumPy oracle on seeded inputs.

Tolerances (BASELINE.json north_star): FP64 fields within 1e-12 relative (max norm) per
step, 1e-9 after 1000 steps; boundary and index handling bit-exact.  The kernels keep the
reference's operation order, so most checks below demand exact equality.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from tests.cases import STEP_CASES, VISC_CASES, build_case, inputs  # noqa: E402

REL_STEP = 1e-12
REL_1000 = 1e-9


def relerr(a, b):
    """max-norm relative error of the density plane and of the mass-flux vector (jx, jy) -- the two
    flux components share one scale (a component that is physically ~0, e.g. jy in an x-driven
    film, has no scale of its own)."""
    a, b = np.asarray(a), np.asarray(b)
    e_rho = np.abs(a[0] - b[0]).max() / max(np.abs(b[0]).max(), 1e-300)
    e_j = np.abs(a[1:] - b[1:]).max() / max(np.abs(b[1:]).max(), 1e-300)
    return float(max(e_rho, e_j))


ARITH = ['exact', 'fast']      # gpf_fv_cfg.arith: bit-identical arithmetic / contracted arithmetic (1e-12 per step)


def make_solver(name, oracle=None, arith='exact'):
    from gapflow_b200.fv import FVSolver
    d, extra = inputs(name)
    o = oracle or build_case(name)
    s = FVSolver(d['grid'], d['geometry'], d['properties'], d['numerics'], topo=o.topo, extra=extra, arith=arith)
    s.set_q(o.q)
    s.pre_run(reset_ekin_old=False)      # dt = CFL*dt_crit of the perturbed field, Ekin_old of _initialize
    return s, o


@pytest.fixture(scope="module")
def gc(golden_dir):
    return np.load(os.path.join(golden_dir, "fv_closures.npz"))


@pytest.fixture(scope="module")
def gs(golden_dir):
    return np.load(os.path.join(golden_dir, "fv_steps.npz"))


def dev(x):
    return torch.as_tensor(np.ascontiguousarray(x), device='cuda')


# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("fn", ["stress_bottom", "stress_top", "stress_avg"])
def test_stress_functions_bit_exact_vs_reference_golden(gc, fn):
    from gapflow_b200 import fv
    q, h = dev(gc['q']), dev(gc['h'])
    for k, (U, V, eta, zeta) in enumerate(gc['visc_params']):
        for ls_name, ls in (('ls0', None), ('ls0', dev(np.zeros_like(gc['Ls']))), ('ls', dev(gc['Ls']))):
            got = getattr(fv, fn)(q, h, float(U), float(V), float(eta), float(zeta), ls).cpu().numpy()
            want = gc[f'{fn}_{k}_{ls_name}']
            assert np.array_equal(got, want), (fn, k, ls_name, relerr(got, want))


def test_eos_functions_vs_reference_golden(gc):
    from gapflow_b200 import fv
    from oracle.make_golden import EOS_PROPS
    # DH, vdW use only + - * / sqrt: exact.  PL, MT, cubic, BWR, Bayada call pow/exp/log: NumPy itself
    # dispatches those to CPU-specific SIMD routines (~1 ulp, e.g. x**3 != correctly rounded x^3 on
    # 5 % of inputs on the build machine), so the reference is only defined to ~1e-15 there.
    exact = {'DH', 'vdW'}
    for prop in EOS_PROPS:
        e = prop['EOS']
        d = dev(gc[f'dens_{e}'])
        p = fv.eos_pressure(d, prop).cpu().numpy()
        c = fv.eos_sound_velocity(d, prop).cpu().numpy()
        wp, wc = gc[f'p_{e}'], gc[f'c_{e}']
        if e in exact:
            assert np.array_equal(p, wp), e
            assert np.array_equal(c, wc, equal_nan=True), e
        else:
            assert np.allclose(p, wp, rtol=1e-13, atol=1e-13 * np.abs(wp).max()), (e, np.abs(p - wp).max())
            assert np.array_equal(np.isnan(c), np.isnan(wc)), e
            m = ~np.isnan(wc)
            assert np.allclose(c[m], wc[m], rtol=1e-13), e


@pytest.mark.parametrize("arith", ARITH)
@pytest.mark.parametrize("name", list(STEP_CASES))
def test_step_vs_reference_golden(gs, name, arith):
    s, o = make_solver(name, arith=arith)
    assert np.array_equal(s.get_q(), gs[f'{name}_q0'])
    sc = s.scalars()
    # BWR: pow/exp -> tolerance (see EOS test); fast arithmetic: tolerance by contract
    # viscosity modifiers: exp / pow / asinh / hypot -> tolerance in both contracts (include/gapflow_b200.h)
    exact = STEP_CASES[name]['yaml'].find('EOS: DH') >= 0 and arith == 'exact' and name not in VISC_CASES
    dt0 = float(gs[f'{name}_dt0'])
    assert sc.dt == dt0 if exact else abs(sc.dt - dt0) <= 1e-12 * dt0
    for nsteps, tag in ((1, '1'), (2, '3')):
        s.step(nsteps)
        q = s.get_q()
        want = gs[f'{name}_q{tag}']
        err = relerr(q, want)
        assert err <= REL_STEP * int(tag), (name, tag, err)      # per-step tolerance x steps taken
        if exact:
            assert np.array_equal(q, want), (name, tag, int((q != want).sum()), err)
        sc = s.scalars()
        dt, res, ekin, vmax, vsound, mass = gs[f'{name}_s{tag}']
        if exact:
            assert sc.dt == dt and sc.v_max == vmax and sc.v_sound == vsound
        else:
            assert abs(sc.dt - dt) <= 1e-12 * dt
        assert abs(sc.ekin - ekin) <= 1e-12 * abs(ekin)          # summation order differs
        assert abs(sc.residual - res) <= 1e-9 * abs(res) + 1e-14  # |Ekin - Ekin_old| cancels digits
        assert abs(s.mass - mass) <= 1e-12 * abs(mass)
        assert sc.step == int(tag) and not sc.stop


@pytest.mark.parametrize("arith", ARITH)
@pytest.mark.parametrize("name", ['journal2d_periodic', 'slider2d_xDN', 'slider2d_allDN_slip'])
def test_multi_tile_grid_vs_oracle(name, arith):
    """Grids spanning several 16x64 tiles with ragged edges; 4 steps against the oracle."""
    import copy
    from gapflow_b200.fv import FVSolver
    from oracle.fv import FVOracle
    from tests.cases import perturb, slip_field
    d, _ = inputs(name)
    d = copy.deepcopy(d)
    g = d['grid']
    scale_x, scale_y = 83 / g['Nx'], 150 / g['Ny']
    g['Nx'], g['Ny'] = 83, 150
    if name == 'slider2d_xDN':
        g['dx'], g['dy'] = g['Lx'] / 83, g['Ly'] / 150
    else:
        g['Lx'], g['Ly'] = g['dx'] * 83, g['dy'] * 150
    extra = slip_field((85, 152), 1e-6) if STEP_CASES[name].get('extra') else None
    o = FVOracle(g, d['geometry'], d['properties'], d['numerics'], extra=extra)
    perturb(o.q, d['properties']['rho0'], max(abs(d['geometry']['U']), abs(d['geometry']['V'])), 4321)
    o.fill_ghosts()
    o.dt = d['numerics']['CFL'] * o.dt_crit
    s = FVSolver(g, d['geometry'], d['properties'], d['numerics'], topo=o.topo, extra=extra, arith=arith)
    s.set_q(o.q)
    s.pre_run(False)
    for k in range(4):
        o.update()
        s.step(1)
        q = s.get_q()
        if arith == 'exact':
            assert np.array_equal(q, o.q), (name, k, int((q != o.q).sum()), relerr(q, o.q))
            assert s.scalars().dt == o.dt
        else:
            assert relerr(q, o.q) <= REL_STEP * (k + 1), (name, k, relerr(q, o.q))
            assert abs(s.scalars().dt - o.dt) <= 1e-12 * o.dt
            # boundary and index handling stay bit-exact: ghost cells are exact functions of the interior
            ref = type(o).__new__(type(o))
            ref.__dict__.update(o.__dict__)
            ref.q = q.copy()
            ref.fill_ghosts()
            assert np.array_equal(ref.q, q)


def test_closure_fields_match_oracle():
    s, o = make_solver('slider2d_allDN_slip')
    o.closures()
    p, wxz, wyz, bulk = s.closures()
    assert np.array_equal(p, o.pressure)
    assert np.array_equal(wxz, o.wall_xz)
    assert np.array_equal(wyz, o.wall_yz)
    assert np.array_equal(bulk, o.bulk)


# ---- viscosity modifiers (stress.py:305-326, viscosity.py) -----------------------------
def _maxrel(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(b).max(), 1e-300))


@pytest.mark.parametrize("arith", ARITH)
@pytest.mark.parametrize("name", VISC_CASES)
def test_viscosity_field_and_closures_vs_oracle(name, arith):
    """The viscosity field the device builds before each pass, and the closure fields evaluated with it,
    against the oracle (== the reference's piezoviscosity / shear_rate_avg / shear_thinning_factor chain,
    tests/test_oracle_vs_reference.py).  Transcendentals: 1e-13 relative, every cell incl. ghosts."""
    s, o = make_solver(name, arith=arith)
    o.closures()
    want = o.viscosity()
    eta = s.viscosity()
    assert eta.shape == want.shape
    assert np.allclose(eta, want, rtol=1e-13, atol=0.0), (name, float(np.abs(eta / want - 1).max()))
    assert np.ptp(want) > 0.0            # the case really has a non-uniform viscosity
    p, wxz, wyz, bulk = s.closures()
    assert _maxrel(p, o.pressure) <= 1e-13
    for got, ref in ((wxz, o.wall_xz), (wyz, o.wall_yz), (bulk, o.bulk)):
        assert np.array_equal(got == 0.0, ref == 0.0)
        assert _maxrel(got, ref) <= 1e-12, (name, _maxrel(got, ref))


@pytest.mark.parametrize("arith", ARITH)
@pytest.mark.parametrize("name", VISC_CASES)
def test_viscosity_cases_ten_steps_vs_oracle(name, arith):
    """10 steps with the viscosity field against the oracle; the stale closure fields afterwards are those
    of the last corrector pass's input (problem.py:532-540), viscosity included."""
    s, o = make_solver(name, arith=arith)
    for _ in range(10):
        o.update()
    s.step(10)
    q = s.get_q()
    assert relerr(q, o.q) <= REL_STEP * 10, (name, relerr(q, o.q))
    assert abs(s.scalars().dt - o.dt) <= 1e-11 * o.dt
    # ghost cells remain exact functions of the interior
    ref = type(o).__new__(type(o))
    ref.__dict__.update(o.__dict__)
    ref.q = q.copy()
    ref.fill_ghosts()
    assert np.array_equal(ref.q, q)
    # o.pressure / o.bulk are the corrector-pass fields of the last step (refreshed at the start of a pass only)
    p, wxz, wyz, bulk = s.closures(stale=True)
    # (the fields of the final q differ from these by ~10 %; 10 steps of 1e-12 in q are amplified by K/p ~ 1e5 in p)
    assert _maxrel(p, o.pressure) <= 1e-6
    assert _maxrel(bulk, o.bulk) <= 1e-6
    assert _maxrel(wxz, o.wall_xz) <= 1e-6


def test_viscosity_multi_tile_ragged_grid_vs_oracle():
    """Barus + Eyring on an 83 x 150 grid (several tiles, ragged edges, D/N x-sides): np.gradient's one-sided
    edge rows and the tile halos see every position; 4 steps."""
    import copy
    from gapflow_b200.fv import FVSolver
    from oracle.fv import FVOracle
    from tests.cases import perturb
    d, _ = inputs('slider2d_barus_eyring')
    d = copy.deepcopy(d)
    g = d['grid']
    g['Nx'], g['Ny'] = 83, 150
    g['dx'], g['dy'] = g['Lx'] / 83, g['Ly'] / 150
    o = FVOracle(g, d['geometry'], d['properties'], d['numerics'])
    perturb(o.q, d['properties']['rho0'], abs(d['geometry']['U']), 4321)
    o.fill_ghosts()
    o.dt = d['numerics']['CFL'] * o.dt_crit
    for arith in ARITH:
        oo = copy.deepcopy(o)
        s = FVSolver(g, d['geometry'], d['properties'], d['numerics'], topo=oo.topo, arith=arith)
        s.set_q(oo.q)
        s.pre_run(False)
        oo.closures()
        assert np.allclose(s.viscosity(), oo.viscosity(), rtol=1e-13, atol=0.0)
        for k in range(4):
            oo.update()
            s.step(1)
            assert relerr(s.get_q(), oo.q) <= REL_STEP * (k + 1), (arith, k, relerr(s.get_q(), oo.q))


def test_viscosity_requires_set_viscosity_and_identity_law():
    """cfg.visc_field without gpf_fv_set_viscosity is a call-sequence error, not a Newtonian run; a block
    with an unknown law name is the identity (viscosity.py:64, 94) and reproduces the Newtonian step."""
    import copy
    import ctypes as C
    from gapflow_b200 import _capi
    from gapflow_b200.fv import FVSolver
    d, _ = inputs('journal2d_periodic')
    o = build_case('journal2d_periodic')
    prop = copy.deepcopy(d['properties'])
    prop['piezo'] = {'name': 'none'}
    s = FVSolver(d['grid'], d['geometry'], prop, d['numerics'], topo=o.topo)
    s.set_q(o.q)
    s.pre_run(False)
    s.step(3)
    s0, _ = make_solver('journal2d_periodic')
    s0.step(3)
    assert np.array_equal(s.get_q(), s0.get_q())
    assert np.array_equal(s.viscosity(), np.full(s.shape, prop['shear']))
    # raw handle without the set-up call
    lib = _capi.load()
    cfg = _capi.make_cfg(d['grid'], d['geometry'], prop, d['numerics'])
    n = C.c_size_t()
    _capi.check(lib.gpf_fv_workspace_bytes(C.byref(cfg), C.byref(n)))
    ws = torch.empty(n.value, dtype=torch.uint8, device='cuda')
    h = C.c_void_p()
    _capi.check(lib.gpf_fv_create(C.byref(cfg), C.c_void_p(ws.data_ptr()), n.value, C.byref(h)))
    assert lib.gpf_fv_step(h, 1, 0, None) == _capi.GPF_E_STATE
    assert b'gpf_fv_set_viscosity' in lib.gpf_last_error()
    lib.gpf_fv_destroy(h)


def test_fill_ghosts_and_host_update():
    s, o = make_solver('slider2d_allDN_slip')
    q = o.q.copy()
    q[:, 0, :] = -1.0
    q[:, -1, :] = -2.0
    q[:, :, 0] = -3.0
    q[:, :, -1] = -4.0
    s.set_q(q)
    s.fill_ghosts()
    assert np.array_equal(s.get_q(), o.q)
    # host-buffer step == device-resident step
    s2, _ = make_solver('slider2d_allDN_slip')
    s2.step(2)
    qh = np.ascontiguousarray(o.q.copy())
    s3, _ = make_solver('slider2d_allDN_slip')
    s3.update_host(qh, 2)
    assert np.array_equal(qh, s2.get_q())


@pytest.mark.parametrize("arith", ARITH)
def test_invalid_state_restores_previous_step_and_stops(arith):
    """problem.py:565-610: NaN / negative density -> q restored, _stop set."""
    s, o = make_solver('journal2d_periodic', arith=arith)
    s.step(1)
    q1 = s.get_q()
    s.set_scalars(dt=1e-3)            # absurd time step -> negative density or NaN
    s.step(1)
    sc = s.scalars()
    assert sc.stop == 1 and (sc.has_nan or sc.has_neg)
    assert sc.step == 1
    assert np.array_equal(s.get_q(), q1)
    s.step(3)                         # further updates are no-ops
    assert np.array_equal(s.get_q(), q1) and s.scalars().step == 1


@pytest.mark.parametrize("name", ['journal2d_periodic', 'slider2d_xDN', 'slider2d_allDN_slip'])
def test_operands_outside_the_fast_range_take_the_general_kernel(name):
    """A pass is a straight-line kernel (quotients over reciprocals, operands in [2^-100, 2^100])
    followed by the general kernel, which redoes the pass with IEEE divisions where a cell left
    that range.  Fluxes of 1e-40 and 1e-300 (subnormal products) force the second; the result must
    stay bit-identical to the NumPy oracle, and a later in-range step must return to the first."""
    s, o = make_solver(name)
    q = o.q.copy()
    q[1, 3, 5] = 1e-40
    q[2, 7, 2] = -3e-300
    q[1, -2, -2] = 7e-35
    o.q[...] = q
    o.fill_ghosts()
    if o.numerics['adaptive']:
        o.dt = o.numerics['CFL'] * o.dt_crit
    s.set_q(o.q)
    s.pre_run(reset_ekin_old=False)
    for _ in range(3):
        o.update()
    s.step(3)
    assert np.array_equal(s.get_q(), o.q)
    sc = s.scalars()
    assert sc.dt == o.dt
    assert abs(sc.residual - o.residual) <= 1e-9 * abs(o.residual)      # the Ekin sum is re-ordered


def test_run_until_converged_freezes_at_the_reference_step():
    """problem.py:401: while not converged and step < max_it."""
    import copy
    from gapflow_b200.fv import FVSolver
    d, _ = inputs('journal1d_periodic')
    d = copy.deepcopy(d)
    d['numerics']['max_it'] = 37
    o = build_case('journal1d_periodic')
    o.max_it = 37
    s = FVSolver(d['grid'], d['geometry'], d['properties'], d['numerics'], topo=o.topo)
    s.set_q(o.q)
    s.pre_run(False)
    o.run()
    s.step(64, until_converged=True)
    sc = s.scalars()
    assert sc.step == o.step == 37
    assert np.array_equal(s.get_q(), o.q)


@pytest.mark.parametrize("arith", ARITH)
@pytest.mark.parametrize("name", ['journal1d_readme', 'journal2d_periodic', 'slider2d_xDN'])
def test_1000_steps(name, arith):
    s, o = make_solver(name, arith=arith)
    n = 1000 if name != 'journal1d_readme' else 200      # README case: max_it = 200
    for _ in range(n):
        o.update()
    s.step(n)
    q = s.get_q()
    err = relerr(q, o.q)
    assert err <= REL_1000, (name, err)
    sc = s.scalars()
    if arith == 'exact':
        assert np.array_equal(q, o.q), (name, int((q != o.q).sum()), err)
        assert sc.dt == o.dt
    assert sc.step == o.step and abs(sc.dt - o.dt) <= 1e-9 * o.dt
    assert abs(sc.simtime - o.simtime) <= 1e-12 * o.simtime


def test_branch_free_division_equals_ieee():
    """closures.cuh::rcp_fast / div_fast against __drcp_rn / __ddiv_rn on 2^31 structured operand pairs."""
    import ctypes as C
    from gapflow_b200 import _capi
    lib = _capi.load()
    out = (C.c_uint64 * 2)()
    _capi.check(lib.gpf_selftest_division(1 << 31, 20251104, out, None))
    assert out[1] == 0, f"{out[1]} quotient mismatches"
    assert out[0] == 0, f"{out[0]} reciprocal mismatches"


# ---- host-buffer step with the copies overlapped chunk by chunk (gpf_fv_update_host, n = 1) ---------------
@pytest.mark.parametrize("arith", ARITH)
@pytest.mark.parametrize("name,nx,ny", [('journal2d_periodic', 150, 70), ('slider2d_xDN', 130, 65),
                                        ('slider2d_allDN_slip', 97, 21), ('journal1d_periodic', 100, 1)])
def test_pipelined_host_step_equals_device_step(name, nx, ny, arith):
    """gpf_fv_update_host(n = 1) moves the field in chunks of tile rows on two copy streams while the pass
    kernels follow chunk by chunk (P(0) P(1) C(0) P(2) C(1) ...): bit-identical to gpf_fv_step, ghost rows
    included, for periodic and Dirichlet / Neumann x sides, ragged last tile rows, and against the oracle."""
    from gapflow_b200.fv import FVSolver
    from oracle.fv import FVOracle
    from tests.cases import perturb, slip_field
    d, _ = inputs(name)
    g = d['grid']
    g['Nx'], g['Ny'] = nx, ny
    g['Lx'], g['Ly'] = g['dx'] * nx, g['dy'] * ny
    extra = slip_field((nx + 2, ny + 2), 1e-6) if STEP_CASES[name].get('extra') else None
    o = FVOracle(g, d['geometry'], d['properties'], d['numerics'], extra=extra)
    perturb(o.q, d['properties']['rho0'], max(abs(d['geometry']['U']), abs(d['geometry']['V'])), 99)
    o.fill_ghosts()
    if d['numerics']['adaptive']:
        o.dt = d['numerics']['CFL'] * o.dt_crit
    assert (nx + 7) // 8 >= 8                     # enough tile rows for the chunked path
    sol = []
    for _ in range(2):
        s = FVSolver(g, d['geometry'], d['properties'], d['numerics'], topo=o.topo, extra=extra, arith=arith)
        s.set_q(o.q)
        s.pre_run(reset_ekin_old=False)
        sol.append(s)
    qh = torch.from_numpy(o.q.copy()).pin_memory().numpy()
    for n in range(3):
        sol[0].step(1)
        sc = sol[1].update_host(qh, 1)
        assert np.array_equal(qh, sol[0].get_q()), (n, int((qh != sol[0].get_q()).sum()))
        s0 = sol[0].scalars()
        assert sc.dt == s0.dt and sc.step == s0.step == n + 1
        # the kinetic-energy sum is folded chunk by chunk: same value up to the summation order
        assert np.isclose(sc.ekin, s0.ekin, rtol=1e-13) and np.isclose(sc.residual, s0.residual, rtol=1e-6, atol=1e-14)
        o.update()
    assert relerr(qh, o.q) <= 10 * REL_STEP          # three steps of the contracted arithmetic; exact: identical
    if arith == 'exact':
        assert np.array_equal(qh, o.q)
    # pageable host memory takes the same path (the copies are then staged by the driver)
    qp = sol[0].get_q().copy()
    sol[0].step(1)
    sol[1].update_host(qp, 1)
    assert np.array_equal(qp, sol[0].get_q())
