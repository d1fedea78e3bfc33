"""GPU parity at BASELINE.json's full size (config 2: 4096 x 4096) through properties that do not
need the oracle on the whole grid (it takes ~70 s per step there):

* locality: with a fixed dt, n steps of a window depend only on the window plus 2n cells around it,
  so the NumPy oracle run on a window + margin must reproduce the GPU's full-grid result inside the
  window bit for bit (exact contract) -- windows across tile corners, at a physical x boundary and
  across the periodic y seam;
* shift invariance: a y-periodic problem whose topography does not depend on y commutes with a
  cyclic shift in y, bit for bit (tiles, halos, index arithmetic), in both arithmetic contracts;
* mass conservation of the fully periodic problem (tests/test_mass_conservation.py:67-77);
* the two arithmetic contracts agree to 1e-12 per step;
* two slabs == one domain, bit for bit.
"""
import copy

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

import bench  # noqa: E402
from tests.test_fv_gpu import relerr  # noqa: E402

N = 4096


def _solver(d, q, topo, arith='exact'):
    from gapflow_b200.fv import FVSolver
    s = FVSolver(d['grid'], d['geometry'], d['properties'], d['numerics'], topo=topo, arith=arith)
    s.set_q(q)
    s.fill_ghosts()
    s.pre_run(reset_ekin_old=False)
    return s


@pytest.fixture(scope="module")
def slider():
    d = bench.slider_inputs(N, N)
    q, topo = bench.initial_fields(d)
    return d, q, topo


def test_config2_windows_match_the_oracle_bit_for_bit(slider):
    from oracle.fv import FVOracle
    d, q, topo = slider
    d = copy.deepcopy(d)
    d['numerics']['adaptive'] = 0
    d['numerics']['dt'] = 3e-10
    s = _solver(d, q, topo)
    q0 = s.get_q()                       # ghosts filled
    steps, m = 3, 8
    s.step(steps)
    sc = s.scalars()
    assert sc.step == steps and not sc.stop and sc.dt == 3e-10
    q1 = s.get_q()
    # windows as (padded row indices, padded column indices), margins included.  A window side that is
    # not a physical boundary of the grid is a fake boundary for the oracle: what it puts there spreads
    # 2 cells per step, so m = 8 cells are discarded on those sides.
    def rng(a, b):
        return list(range(a, b))
    windows = [(rng(2040 - m, 2120 + m), rng(2030 - m, 2130 + m)),           # tile corners in the middle
               (rng(0, 60 + m), rng(1000 - m, 1090 + m)),                    # x-low physical boundary (ghost row 0)
               (rng(N + 2 - 60 - m, N + 2), rng(3000 - m, 3090 + m)),        # x-high physical boundary
               (rng(511 - m, 551 + m), rng(N + 1 - 35 - m, N + 1) + rng(1, 36 + m))]   # across the periodic y seam
    for rows, cols in windows:
        ix = np.ix_(range(3), rows, cols)
        g = dict(d['grid'])
        g['Nx'], g['Ny'] = len(rows) - 2, len(cols) - 2
        g['Lx'], g['Ly'] = g['dx'] * g['Nx'], g['dy'] * g['Ny']
        o = FVOracle(g, d['geometry'], d['properties'], d['numerics'], topo=topo[np.ix_(range(topo.shape[0]), rows, cols)].copy())
        o.q[...] = q0[ix]
        o.dt = 3e-10
        for _ in range(steps):
            o.update()
        assert o.dt == 3e-10
        lo = 0 if rows[0] == 0 else m
        hi = len(rows) if rows[-1] == N + 1 else len(rows) - m
        inner = (slice(None), slice(lo, hi), slice(m, len(cols) - m))
        got, ref = q1[ix][inner], o.q[inner]
        assert np.array_equal(got, ref), (rows[0], cols[0], int((got != ref).sum()))


@pytest.mark.parametrize("arith", ['exact', 'fast'])
def test_shift_invariance_in_y_full_size(slider, arith):
    d, q, topo = slider
    k = 1000
    s1 = _solver(d, q, topo, arith)
    qs = q.copy()
    qs[:, :, 1:-1] = np.roll(q[:, :, 1:-1], k, axis=2)
    s2 = _solver(d, qs, topo, arith)
    s1.step(5)
    s2.step(5)
    a, b = s1.get_q(), s2.get_q()
    assert np.array_equal(np.roll(a[:, :, 1:-1], k, axis=2), b[:, :, 1:-1])
    assert s1.scalars().dt == s2.scalars().dt


def test_mass_conservation_periodic_full_size():
    from gapflow_b200 import config
    from gapflow_b200.topography import make_topography
    y = ("options:\n    output: data/journal\n    write_freq: 1000\n    silent: True\n"
         f"grid:\n    dx: 1.e-5\n    dy: 1.e-5\n    Nx: {N}\n    Ny: {N}\n"
         "    xE: ['P', 'P', 'P']\n    xW: ['P', 'P', 'P']\n    yS: ['P', 'P', 'P']\n    yN: ['P', 'P', 'P']\n"
         "geometry:\n    type: journal\n    CR: 1.e-2\n    eps: 0.7\n    U: 0.1\n    V: 0.\n"
         "numerics:\n    CFL: 0.5\n    adaptive: 1\n    tol: 1e-8\n    dt: 1e-10\n    max_it: 5000\n"
         "properties:\n    shear: 0.0794\n    bulk: 0.\n    EOS: DH\n    P0: 101325.\n    rho0: 877.7007\n"
         "    T0: 323.15\n    C1: 3.5e10\n    C2: 1.23\n")
    d = config.read_yaml_string(y)
    q, topo = bench.initial_fields(d, seed=5)
    s = _solver(d, q, topo)
    m0 = s.mass
    s.step(10)
    sc = s.scalars()
    assert sc.step == 10 and not sc.stop
    assert abs(s.mass / m0 - 1.0) <= 1e-11


def test_contracts_agree_full_size(slider):
    d, q, topo = slider
    a, b = _solver(d, q, topo, 'exact'), _solver(d, q, topo, 'fast')
    a.step(4)
    b.step(4)
    assert relerr(b.get_q(), a.get_q()) <= 4e-12
    assert abs(b.scalars().dt - a.scalars().dt) <= 1e-12 * a.scalars().dt


def test_two_slabs_equal_one_domain_full_size(slider):
    from gapflow_b200.dist import InProcessSlabs
    d, q, topo = slider
    single = _solver(d, q, topo)
    q0 = single.get_q()
    slabs = InProcessSlabs(d, q0, topo, 2)
    single.step(3)
    slabs.step(3)
    assert np.array_equal(slabs.gather_q(), single.get_q())
    for sc in slabs.scalars():
        assert sc.dt == single.scalars().dt


def test_config4_gp_inference_vs_oracle_on_a_subsample():
    """BASELINE config 4: N = 1026^2 query cells x M = 10^4 training points.  Mean over the whole
    grid, mean + variance over a 16 384-cell block; the SciPy oracle checks a 2 048-cell random
    subsample of each (the full triangular solve is ~10^14 flops on the CPU); 1e-9 relative."""
    from gapflow_b200.gp import GPModel
    from oracle import gp as G
    from tests.gp_cases import query_set, training_set
    M, Nq = 10000, 1026 * 1026
    c = training_set(M, 'shear_x')
    gp = GPModel(M, len(c['dims']), c['n_out'])
    gp.factor(c['X'], c['Y'], c['log_amp'], c['log_scale'], c['Yerr'])
    xt = query_set(Nq, c)
    xd = torch.as_tensor(xt, device='cuda')
    mean = gp.repredict_mean(xd).cpu().numpy()
    blk = slice(500000, 500000 + 16384)
    m2, var = gp.repredict_mean_var(xd[blk].contiguous(), 1.5e-8)
    m2, var = m2.cpu().numpy(), var.cpu().numpy()
    assert np.array_equal(m2, mean[:, blk])                      # same kernel arithmetic, any batch
    L = G.build(c['X'], c['Yerr'], c['log_amp'], c['log_scale'])
    a = G.alpha(L, c['Y'])
    rng = np.random.default_rng(3)
    pick = np.sort(rng.choice(Nq, 2048, replace=False))
    mref = G.repredict_mean(c['X'], c['log_amp'], c['log_scale'], a, xt[pick])
    assert np.abs(mean[:, pick] - mref).max() <= 1e-9 * np.abs(mref).max()
    sub = np.sort(rng.choice(16384, 2048, replace=False))
    _, vref = G.repredict_mean_var(c['X'], c['log_amp'], c['log_scale'], L, a, 1.5e-8, xt[blk][sub])
    assert np.abs(var[sub] - vref).max() <= 1e-9 * np.abs(vref).max()
    assert gp.argmax == int(np.argmax(var)) and abs(gp._vmax_norm - var.max()) == 0.0


# ---- BASELINE config 5 / north_star target grid: the periodic journal at 8192 x 8192 -----------------------
N5 = 8192


@pytest.fixture(scope="module")
def journal8192():
    d = bench.journal_inputs(N5, N5)
    q, topo = bench.initial_fields(d)
    # exactly periodic gap field: the generator's ghost rows (x = -dx/2, Lx + dx/2) equal the periodic images only
    # to rounding, and the seam tests below compare bit for bit with runs that see the image itself
    topo[:, 0], topo[:, -1] = topo[:, -2], topo[:, 1]
    topo[:, :, 0], topo[:, :, -1] = topo[:, :, -2], topo[:, :, 1]
    return d, q, topo


def test_config5_windows_match_the_oracle_bit_for_bit(journal8192):
    """Locality at 8192^2, fixed dt, exact contract: windows across tile corners, across the periodic x seam
    (rows Nx-k .. Nx, 1 .. k: what the ring exchange carries on slabs), across the periodic y seam and across
    both seams at once (the corner ghosts)."""
    from oracle.fv import FVOracle
    d, q, topo = journal8192
    d = copy.deepcopy(d)
    d['numerics']['adaptive'] = 0
    d['numerics']['dt'] = 2e-10
    s = _solver(d, q, topo)
    q0 = s.get_q()
    steps, m = 3, 8
    s.step(steps)
    sc = s.scalars()
    assert sc.step == steps and not sc.stop and sc.dt == 2e-10
    q1 = s.get_q()
    del s
    torch.cuda.empty_cache()

    def rng(a, b):
        return list(range(a, b))
    seam_x = rng(N5 + 1 - 30 - m, N5 + 1) + rng(1, 31 + m)       # interior rows Nx-37 .. Nx, 1 .. 38
    seam_y = rng(N5 + 1 - 40 - m, N5 + 1) + rng(1, 41 + m)
    windows = [(rng(4090 - m, 4170 + m), rng(4060 - m, 4200 + m)),          # tile corners in the middle
               (seam_x, rng(5000 - m, 5100 + m)),                            # periodic x seam
               (rng(6001 - m, 6041 + m), seam_y),                            # periodic y seam
               (seam_x, seam_y)]                                             # both: the corner ghost cells
    for rows, cols in windows:
        ix = np.ix_(range(3), rows, cols)
        g = dict(d['grid'])
        g['Nx'], g['Ny'] = len(rows) - 2, len(cols) - 2
        g['Lx'], g['Ly'] = g['dx'] * g['Nx'], g['dy'] * g['Ny']
        o = FVOracle(g, d['geometry'], d['properties'], d['numerics'], topo=topo[np.ix_(range(topo.shape[0]), rows, cols)].copy())
        o.q[...] = q0[ix]
        o.dt = 2e-10
        for _ in range(steps):
            o.update()
        inner = (slice(None), slice(m, len(rows) - m), slice(m, len(cols) - m))
        got, ref = q1[ix][inner], o.q[inner]
        assert np.array_equal(got, ref), (rows[0], cols[0], int((got != ref).sum()))
    # the ghost layers of the full field are the periodic images of the interior (problem.py:681-707)
    assert np.array_equal(q1[:, 0, 1:-1], q1[:, N5, 1:-1]) and np.array_equal(q1[:, N5 + 1, 1:-1], q1[:, 1, 1:-1])
    assert np.array_equal(q1[:, :, 0], q1[:, :, N5]) and np.array_equal(q1[:, :, N5 + 1], q1[:, :, 1])


@pytest.mark.parametrize("arith", ['exact', 'fast'])
def test_config5_shift_invariance_in_x_and_y(journal8192, arith):
    """The fully periodic problem commutes with cyclic shifts of the interior (solution AND gap field) in x and
    in y, bit for bit: tiles, halos, the persistent-CTA walk and the seam index arithmetic at 8192^2."""
    d, q, t1 = journal8192
    kx, ky = 3001, 1234
    s1 = _solver(d, q, t1, arith)
    s1.step(4)
    a = s1.get_q()
    dt1 = s1.scalars().dt
    del s1
    torch.cuda.empty_cache()

    def shifted(f):
        out = f.copy()
        out[:, 1:-1, 1:-1] = np.roll(f[:, 1:-1, 1:-1], (kx, ky), axis=(1, 2))
        out[:, 0], out[:, -1] = out[:, -2], out[:, 1]
        out[:, :, 0], out[:, :, -1] = out[:, :, -2], out[:, :, 1]
        return out
    s2 = _solver(d, shifted(q), shifted(t1), arith)
    s2.step(4)
    b = s2.get_q()
    assert s2.scalars().dt == dt1
    assert np.array_equal(np.roll(a[:, 1:-1, 1:-1], (kx, ky), axis=(1, 2)), b[:, 1:-1, 1:-1])
