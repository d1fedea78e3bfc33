"""GPU parity tests of the GP-coupled step (BASELINE configs 3 and 4 scaled down): the surrogate variant of
the pass kernels, the device mock MD, and the whole Problem.update() with active learning, against
oracle/surrogate.py on the same seeded inputs.

Tolerances (BASELINE.json north_star): GP mean / variance 1e-9 relative, FV fields 1e-12 per step given
the same closure inputs.  Hyperparameters, mock-MD draws and the initial design are explicit inputs,
identical on both sides (oracle/surrogate.py header)."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import surrogate as OS  # noqa: E402
from tests import surrogate_cases as SC  # noqa: E402
from tests.test_fv_gpu import relerr  # noqa: E402

REL_GP = 1e-9


def build(name, optimizer=SC.fixed_optimizer, arith='exact', al=True, **kw):
    """(inputs, oracle, product solver, product driver) primed with the same perturbed field."""
    from gapflow_b200.fv import FVSolver
    from gapflow_b200.surrogate import Database, Mock, SurrogateDriver
    d, _ = SC.inputs(name, **kw)
    g = d['grid']
    if not al:
        for k in ('press', 'shear'):
            d['gp'][k]['active_learning'] = False
    extra = SC.slip_length(g['Nx'], g['Ny'])
    odb = OS.Database(OS.MockMD(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS), d['db'])
    o = OS.SurrogateFVOracle(g, d['geometry'], d['properties'], d['numerics'], d['gp'], odb, extra=extra,
                             write_freq=d['options']['write_freq'], optimizer=optimizer)
    o.q[...] = SC.perturb(o.q)
    o.fill_ghosts()
    s = FVSolver(g, d['geometry'], d['properties'], d['numerics'], topo=o.topo, extra=extra, arith=arith)
    s.set_q(o.q)
    pdb = Database(Mock(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS), d['db'])
    drv = SurrogateDriver(s, d['gp'], pdb, g['dim'], write_freq=d['options']['write_freq'], optimizer=optimizer)
    return d, o, s, drv


def test_device_mock_md_vs_oracle():
    """md/mock.py:81-107 on the device (gpf_stress_bottom / gpf_stress_top / gpf_eos_pressure) for a batch."""
    from gapflow_b200.surrogate import Mock
    d, _ = SC.inputs('slip2d', nx=12, ny=10)
    rng = np.random.default_rng(3)
    X = np.column_stack([rng.uniform(0.7, 0.9, 64), rng.uniform(0.02, 0.08, 64), rng.uniform(-0.01, 0.01, 64),
                         rng.uniform(10., 14., 64), rng.uniform(-0.01, 0.01, 64), rng.uniform(-0.01, 0.01, 64),
                         rng.uniform(0., 1., 64)])
    Y, Ye = Mock(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS).run_batch(X)
    md = OS.MockMD(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS)
    for k in range(len(X)):
        Yo, Yeo = md.run(X[k], k)
        assert np.allclose(Y[k], Yo, rtol=1e-12, atol=1e-15), (k, np.abs(Y[k] - Yo).max())
        assert np.array_equal(Ye[k], Yeo)


@pytest.mark.parametrize('arith', ['exact', 'fast'])
@pytest.mark.parametrize('name,kw', [('slip1d', dict(nx=100)), ('slip2d', dict(nx=70, ny=37))])
def test_surrogate_pass_kernels_with_given_means(name, kw, arith):
    """The SG variant in isolation: arbitrary smooth arrays stand in for the GP means of both passes (the
    same arrays are written into the oracle's closure fields), so the comparison sees only the kernel:
    pressure of tile + halo cells (ghost cells and ragged tiles included), own wall shear of the interior."""
    from gapflow_b200 import _capi
    d, o, s, drv = build(name, arith=arith, **kw)
    lib = _capi.load()
    shape = o.q.shape[1:]
    Np = shape[0] * shape[1]
    rng = np.random.default_rng(11)
    i, j = np.meshgrid(np.arange(shape[0]), np.arange(shape[1]), indexing='ij')
    means = []
    for r in range(2):                                   # predictor, corrector
        p = 1.0 + 0.1 * np.sin(0.3 * i + r) * np.cos(0.2 * j) + 0.01 * rng.standard_normal(shape)
        xz = 1e-2 * np.stack([np.cos(0.1 * i + 0.3 * j + r), -np.sin(0.2 * i - 0.1 * j)])
        yz = 1e-3 * np.stack([np.sin(0.15 * i + r), np.cos(0.25 * j)])
        means.append((p, xz, yz))
    scales = (3.7, 0.9, 1.3)
    s.pre_run(False)
    o.dt = s.scalars().dt
    # oracle: the update loop of problem.py:524-563 with the given closure fields
    from oracle.fv import predictor_corrector, source
    q0 = o.q.copy()
    for r, dd in enumerate(o.directions()):
        super(OS.SurrogateFVOracle, o).closures()
        p, xz, yz = means[r]
        o.pressure[...] = p * scales[0]
        o.wall_xz[4], o.wall_xz[10] = xz[0] * scales[1], xz[1] * scales[1]
        if name == 'slip2d':
            o.wall_yz[3], o.wall_yz[9] = yz[0] * scales[2], yz[1] * scales[2]
        fX, fY = predictor_corrector(o.q, o.pressure, o.bulk, dd)
        src = source(o.q, o.topo, o.bulk, o.wall_xz[:6] + o.wall_yz[:6], o.wall_xz[6:] + o.wall_yz[6:])
        o.q[...] = o.q - o.dt * (fX / o.grid['dx'] + fY / o.grid['dy'] - src)
        o.fill_ghosts()
    o.q[...] = (o.q + q0) / 2.0
    # product
    keep = []
    for r in range(2):
        p, xz, yz = (torch.as_tensor(np.ascontiguousarray(a), device='cuda') for a in means[r])
        keep += [p, xz, yz]
        _capi.check(lib.gpf_fv_set_surrogate(s.h, C.c_void_p(p.data_ptr()), scales[0], C.c_void_p(xz.data_ptr()), scales[1],
                                             C.c_void_p(yz.data_ptr()) if name == 'slip2d' else None, scales[2]))
        _capi.check(lib.gpf_fv_pass(s.h, r, s._stream()))
    rec = torch.zeros(5, dtype=torch.float64, device='cuda')
    _capi.check(lib.gpf_fv_reduce_local(s.h, 0, C.c_void_p(rec.data_ptr()), s._stream()))
    c2 = 2.5
    _capi.check(lib.gpf_fv_finish_surrogate(s.h, _capi.FIN_STEP, C.c_void_p(rec.data_ptr()), 1, c2, s._stream()))
    q = s.get_q()
    err = relerr(q, o.q)
    assert err <= 1e-12, err
    if arith == 'exact':
        assert np.array_equal(q, o.q)                    # plain IEEE divisions, -fmad=false: the oracle's roundings
    sc = s.scalars()
    assert sc.step == 1 and np.isclose(sc.v_sound, np.sqrt(c2), rtol=1e-15)
    vmax = np.sqrt((o.q[1]**2 + o.q[2]**2) / o.q[0]).max()
    assert np.isclose(sc.dt, d['numerics']['CFL'] * min(o.grid['dx'], o.grid['dy']) / (vmax + np.sqrt(c2)), rtol=1e-13)
    # gpf_fv_step is refused while surrogate arrays are set, and works again after they are cleared
    assert lib.gpf_fv_step(s.h, 1, 0, s._stream()) == _capi.GPF_E_INVALID
    assert b'surrogate' in lib.gpf_last_error()
    _capi.check(lib.gpf_fv_set_surrogate(s.h, None, 1.0, None, 1.0, None, 1.0))
    s.step(1)
    assert s.scalars().step == 2


def _compare_models(o, drv, rtol=REL_GP):
    for k, mo in o.models.items():
        mp = drv.models[k]
        assert mp.last_fit_train_size == mo.last_fit_train_size and mp.n_train == mo.n_train, k
        assert mp._pause == mo._pause and mp._step == mo._step, k
        assert np.isclose(mp.maximum_variance, mo.maximum_variance, rtol=1e-6, atol=1e-14 * mo.Yscale**2), k
        assert np.isclose(mp.variance_tol, mo.variance_tol, rtol=1e-12), k


def test_device_mock_md_and_normalisers_stay_on_the_device():
    """Mock.run_batch and the max-abs normalisers on CUDA tensors: same numbers as the host route."""
    from gapflow_b200.surrogate import Mock, _normalizer
    d, _ = SC.inputs('slip2d', nx=12, ny=10)
    rng = np.random.default_rng(4)
    X = np.column_stack([rng.uniform(0.7, 0.9, 500), rng.uniform(0.02, 0.08, 500), rng.uniform(-0.01, 0.01, 500),
                         rng.uniform(10., 14., 500), rng.uniform(-0.01, 0.01, 500), rng.uniform(-0.01, 0.01, 500),
                         rng.uniform(0., 1., 500)])
    md = Mock(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS)
    Yh, Eh = md.run_batch(X)
    Yd, Ed = md.run_batch(torch.as_tensor(X, device='cuda'))
    assert Yd.is_cuda and Ed.is_cuda
    assert np.array_equal(Yd.cpu().numpy(), Yh) and np.array_equal(Ed.cpu().numpy(), Eh)
    assert np.array_equal(_normalizer(Yd).cpu().numpy(), _normalizer(Yh))
    assert np.array_equal(_normalizer(torch.as_tensor(X, device='cuda')).cpu().numpy(), _normalizer(X))


@pytest.mark.parametrize('name,kw,steps', [
    ('slip1d', dict(nx=100, atol=0.6, al_steps=3), 8),           # BASELINE config 3 scaled: active learning fires
    ('slip2d', dict(nx=24, ny=20, atol=0.8, al_steps=2), 4),     # BASELINE config 4 scaled: press + xz + yz models
    ('slip1d', dict(nx=64, wf=3), 4),                            # variance only before output (problem.py:530)
])
def test_gp_coupled_steps_vs_oracle(name, kw, steps):
    """Problem._pre_run + n x Problem.update() with GP closures: database, trainings, fields and loop scalars."""
    al = 'wf' not in kw
    d, o, s, drv = build(name, al=al, **kw)
    o.pre_run()
    drv.pre_run()
    assert drv.database.size == o.database.size == d['db']['init_size']
    assert np.allclose(drv.database._Ytrain, o.database._Ytrain, rtol=1e-12, atol=1e-15)
    sc = s.scalars()
    assert np.isclose(sc.dt, o.dt, rtol=REL_GP) and np.isclose(sc.v_sound, o.v_sound, rtol=REL_GP)
    for k, mo in o.models.items():
        mf = drv.models[k].mean_field().cpu().numpy()
        want = mo._infer(compute_var=False)[0]
        assert np.abs(mf - want).max() <= REL_GP * np.abs(want).max(), k
        v = drv.models[k].variance.cpu().numpy()
        assert np.allclose(v, mo._predictive_var, rtol=1e-6, atol=1e-9 * np.exp(mo.log_amp) * mo.Yscale**2), k
    for n in range(steps):
        o.update()
        drv.update(n)
        assert drv.database.size == o.database.size, n
        assert np.allclose(drv.database._Xtrain, o.database._Xtrain, rtol=1e-8, atol=1e-12), n
        sc = s.scalars()
        assert sc.step == o.step == n + 1
        err = relerr(s.get_q(), o.q)
        assert err <= 1e-8, (n, err)
        assert np.isclose(sc.dt, o.dt, rtol=1e-8) and np.isclose(sc.residual, o.residual, rtol=1e-5, atol=1e-12), n
        _compare_models(o, drv)
    if al:
        assert o.database.size > d['db']['init_size']           # the case really exercised active learning


def test_default_bfgs_training_close_to_oracle_bfgs():
    """_train with the default optimiser on both sides (SciPy BFGS; objective and gradient from the device here,
    from NumPy in the oracle): same optimum within the optimiser's own tolerance, fields close after steps."""
    d, o, s, drv = build('slip1d', optimizer=None, al=False, nx=64, init=12)
    o.pre_run()
    drv.pre_run()
    for k, mo in o.models.items():
        mp = drv.models[k]
        assert np.isclose(mp.gp.log_amp, mo.log_amp, rtol=1e-3, atol=1e-3), k
        assert np.allclose(mp.gp.log_scale, mo.log_scale, rtol=1e-3, atol=1e-3), k
    for n in range(3):
        o.update()
        drv.update(n)
    assert relerr(s.get_q(), o.q) <= 1e-5


def test_problem_with_gp_from_yaml_runs_and_writes_outputs(tmp_path):
    """Problem.from_string with gp + db sections (problem.py:207-249 builds the Database around the mock MD):
    run() to max_it, outputs sol.nc / history.csv / gp_zz.csv / gp_xz.csv / gp_*.txt / Xtrain.npy."""
    from gapflow_b200.problem import Problem
    from gapflow_b200.surrogate import Database, Mock
    out = str(tmp_path / 'slip')
    d, text = SC.inputs('slip1d', nx=64, max_it=12, wf=5, out=out, silent=False, atol=0.8, al_steps=2)
    # as examples/slip_1d_lj_mock.py:94-107 builds it: runner, database, extra field
    database = Database(Mock(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS), d['db'])
    problem = Problem(d['options'], d['grid'], d['numerics'], d['properties'], d['geometry'], d['gp'], database,
                      extra_field=SC.slip_length(64, 1))
    problem.run()
    assert problem.step == 12 and np.isfinite(problem.q).all()
    assert set(problem.gp_models) == {'zz', 'xz'}
    p = problem.pressure.pressure
    assert np.array_equal(p, problem.gp_models['zz'].mean_field().cpu().numpy())
    w = problem.wall_stress_xz.full
    mf = problem.gp_models['xz'].mean_field().cpu().numpy()
    assert np.array_equal(w[4], mf[0]) and np.array_equal(w[10], mf[1])
    files = set(os.listdir(problem.outdir))
    assert {'sol.nc', 'history.csv', 'gp_zz.csv', 'gp_xz.csv', 'gp_zz.txt', 'gp_xz.txt', 'Xtrain.npy',
            'Ytrain.npy', 'Ytrain_err.npy', 'config.yml'} <= files, files
    rows = open(os.path.join(problem.outdir, 'gp_zz.csv')).read().strip().splitlines()
    assert rows[0].split(',')[:3] == ['step', 'database_size', 'variance'] and len(rows) >= 2
    assert problem.database.size >= 10
    # host edit of q between steps: dt_crit / cfl / v_sound of the edited field take the GP sound speed (stress.py:522-537),
    # as the reference's properties would (problem.py:349-358), not the fixed-form law
    sc0 = problem._solver.scalars()
    problem.q[0] *= 1.0 + 1e-3
    vs = problem.pressure.v_sound
    zz = problem.gp_models['zz']
    want = np.sqrt(zz.v_sound2(0))
    assert np.isclose(vs, want, rtol=1e-12) and vs != sc0.v_sound
    assert np.isclose(problem.dt_crit, min(d['grid']['dx'], d['grid']['dy']) / (problem.v_max + vs), rtol=1e-12)
    # the YAML-only route (no slip-length field: the shear model takes density and flux)
    _, text2 = SC.inputs('slip1d', nx=48, max_it=4, wf=2, out=str(tmp_path / 'yaml'), silent=True, sx='[0, 1]')
    p2 = Problem.from_string(text2)
    p2.run()
    assert p2.step == 4 and np.isfinite(p2.q).all() and p2.database.size >= 10


def test_surrogate_refusals():
    """viscosity-field solvers and slab solvers have no surrogate variant: loud errors, not silent fixed-form."""
    from gapflow_b200 import _capi
    from gapflow_b200.fv import FVSolver
    from tests.cases import VISC_CASES, build_case, inputs
    lib = _capi.load()
    name = VISC_CASES[0]
    d, extra = inputs(name)
    o = build_case(name)
    s = FVSolver(d['grid'], d['geometry'], d['properties'], d['numerics'], topo=o.topo, extra=extra)
    buf = torch.zeros(o.q.shape[1] * o.q.shape[2], dtype=torch.float64, device='cuda')
    rc = lib.gpf_fv_set_surrogate(s.h, C.c_void_p(buf.data_ptr()), 1.0, None, 1.0, None, 1.0)
    assert rc == _capi.GPF_E_INVALID and b'viscosity' in lib.gpf_last_error()


# ---- slabs: training set replicated, query cells partitioned (BASELINE.json north_star) ---------------------
def _run_thread_slabs(d, q_global, topo, extra, world, steps, optimizer=SC.fixed_optimizer):
    """all slabs on one GPU, one thread per rank (gapflow_b200.surrogate.ThreadColl stands in for
    torch.distributed); returns per-rank (slab q, Xtrain, scalars, models)."""
    import threading
    from gapflow_b200 import _capi
    from gapflow_b200.dist import SlabSolver, slab_rows, x_periodic
    from gapflow_b200.surrogate import Database, Mock, SlabSurrogateDriver, ThreadColl
    tc = ThreadColl(world)
    rows = slab_rows(d['grid']['Nx'], world)
    out, errs = [None] * world, []

    def work(r):
        try:
            torch.cuda.set_device(0)
            sl = SlabSolver.from_global(d, None, topo, extra, rank=r, world=world,
                                        exchanger=tc.exchanger(r, x_periodic(d['grid'])))
            sl._fold(1, _capi.FIN_PRERUN)                # kinetic_energy_old of the _initialize state
            start, n = rows[r]
            sl.set_q(np.ascontiguousarray(q_global[:, start:start + n + 2]))
            db = Database(Mock(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS), d['db'])
            drv = SlabSurrogateDriver(sl, d['gp'], db, d['grid']['dim'], tc.view(r),
                                      write_freq=d['options']['write_freq'], optimizer=optimizer)
            drv.pre_run()
            for k in range(steps):
                drv.update(k)
            out[r] = (sl.get_q(), db._Xtrain.copy(), sl.fv.scalars(), drv)
        except BaseException as e:                       # release the other ranks instead of hanging them
            errs.append((r, repr(e)))
            tc._barrier.abort()

    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join(300)
    assert not errs, errs
    return out, rows


@pytest.mark.parametrize('name,kw,world,steps', [
    ('slip2d', dict(nx=24, ny=20, atol=0.8, al_steps=2), 2, 4),
    ('slip1d', dict(nx=100, atol=0.6, al_steps=3), 3, 6),
])
def test_slab_coupled_steps_match_single_domain_and_oracle(name, kw, world, steps):
    """N slabs == one domain bit for bit (same cells, same arithmetic, folds in rank order), and both within the
    GP tolerance of the oracle; the database grows by the same samples on every rank."""
    d, o, s, drv = build(name, **kw)
    q0 = o.q.copy()
    o.pre_run()
    drv.pre_run()
    for k in range(steps):
        o.update()
        drv.update(k)
    q1 = s.get_q()
    out, rows = _run_thread_slabs(d, q0, o.topo, o.extra, world, steps)
    for r, (start, n) in enumerate(rows):
        q, X, sc, sd = out[r]
        assert np.array_equal(q[:, 1:-1], q1[:, start + 1:start + n + 1]), (r, relerr(q[:, 1:-1], q1[:, start + 1:start + n + 1]))
        assert np.array_equal(X, drv.database._Xtrain), r
        s1 = s.scalars()
        assert sc.dt == s1.dt and sc.step == s1.step == steps, r
        assert np.isclose(sc.residual, s1.residual, rtol=1e-10), r      # the kinetic-energy sum is folded rank by rank
        for k, m in sd.models.items():
            assert m.n_train == drv.models[k].n_train and m.argmax == drv.models[k].argmax, (r, k)
            assert m.maximum_variance == drv.models[k].maximum_variance, (r, k)
    assert drv.database.size == o.database.size > d['db']['init_size']
    assert relerr(q1, o.q) <= 1e-8


NCCL_WORKER = r"""
import os, sys
sys.path.insert(0, sys.argv[1])
import numpy as np, torch, torch.distributed as dist
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
rank, world = dist.get_rank(), dist.get_world_size()
from oracle import surrogate as OS
from tests import surrogate_cases as SC
from tests.test_fv_gpu import relerr
from gapflow_b200.dist import SlabSolver, slab_rows
from gapflow_b200.surrogate import Database, Mock, SlabSurrogateDriver, DistColl
d, _ = SC.inputs('slip2d', nx=48, ny=20, atol=0.8, al_steps=2)
g = d['grid']
extra = SC.slip_length(g['Nx'], g['Ny'])
odb = OS.Database(OS.MockMD(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS), d['db'])
o = OS.SurrogateFVOracle(g, d['geometry'], d['properties'], d['numerics'], d['gp'], odb, extra=extra,
                         write_freq=d['options']['write_freq'], optimizer=SC.fixed_optimizer)
o.q[...] = SC.perturb(o.q)
o.fill_ghosts()
sl = SlabSolver.from_global(d, o.q, o.topo, extra)
db = Database(Mock(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS), d['db'])
drv = SlabSurrogateDriver(sl, d['gp'], db, g['dim'], DistColl(), write_freq=d['options']['write_freq'],
                          optimizer=SC.fixed_optimizer)
o.pre_run()
drv.pre_run()
for k in range(4):
    o.update()
    drv.update(k)
q = sl.gather_q()
sc = sl.fv.scalars()
assert db.size == odb.size > d['db']['init_size'], (db.size, odb.size)
assert np.allclose(db._Xtrain, odb._Xtrain, rtol=1e-8, atol=1e-12)
assert relerr(q, o.q) <= 1e-8, relerr(q, o.q)
assert sc.step == o.step and abs(sc.dt - o.dt) <= 1e-8 * o.dt
dist.barrier()
if rank == 0:
    print("NCCL_SURROGATE_OK")
dist.destroy_process_group()
"""


def test_nccl_slab_coupled_steps_vs_oracle(tmp_path):
    import socket
    import subprocess
    import sys
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(n, 4)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "worker.py"
    script.write_text(NCCL_WORKER)
    sock = socket.socket()
    sock.bind(("127.0.0.1", 0))
    port = sock.getsockname()[1]
    sock.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script), root],
                       capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0 and "NCCL_SURROGATE_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("world", [2])
def test_problem_with_gp_on_slabs_writes_the_same_files(tmp_path, world):
    """Problem.run() with GP closures on x-slabs (ranks = threads on one GPU): same sol.nc (variance fields
    included), same training database and hyper-parameter history as the single-domain run."""
    from gapflow_b200.problem import Problem
    from gapflow_b200.surrogate import Database, Mock
    from tests.test_problem_gpu import _read_nc, _run_problem_threads

    def make(par, tag):
        d, _ = SC.inputs('slip1d', nx=64, max_it=10, wf=5, out=str(tmp_path / tag), silent=False, atol=0.8, al_steps=2)
        database = Database(Mock(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS), d['db'])
        return Problem(d['options'], d['grid'], d['numerics'], d['properties'], d['geometry'], d['gp'], database,
                       extra_field=SC.slip_length(64, 1), parallel=par)
    one = make(None, 'one')
    one.run()
    ranks = _run_problem_threads(lambda par: make(par, 'slabs'), world,
                                 collect=lambda p: (p.outdir, p.step, p.database.size, p.database._Xtrain.copy(), p.q.copy()))
    a, b = _read_nc(os.path.join(one.outdir, 'sol.nc')), _read_nc(os.path.join(ranks[0][0], 'sol.nc'))
    assert {'pressure_var', 'wall_stress_xz_var'} <= set(a) and sorted(a) == sorted(b)
    assert a['pressure_var'].shape == a['pressure'].shape and np.isfinite(a['pressure_var']).all()
    assert a['pressure_var'].max() > 0.0                       # frames after the first step hold a computed variance (gp.py:468);
    #                                                            frame 0 the zeros the reference's field starts with
    for k in a:
        assert np.array_equal(a[k], b[k]), k
    for _, step, size, X, q in ranks:
        assert step == one.step and size == one.database.size
        assert np.array_equal(X, one.database._Xtrain)
        assert np.array_equal(q, one.q)
    for name in ('gp_zz.csv', 'gp_xz.csv', 'Xtrain.npy'):
        assert open(os.path.join(one.outdir, name), 'rb').read() == open(os.path.join(ranks[0][0], name), 'rb').read(), name
