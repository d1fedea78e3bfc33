"""GPU tests of the slab decomposition (gapflow_b200/dist.py).

* one GPU: all ranks' slabs in one process, driven in lock step (InProcessSlabs) -- every
  device-side piece of the multi-GPU path (HALO sides, pack/unpack, per-rank records, gathered
  finish, ring-edge ghost rows) against the single-domain solver and the oracle;
* >= 2 GPUs: the real thing, one process per GPU over NCCL (skipped on a 1-GPU box).

q and dt must be bit-identical to the single-domain run for any number of ranks; the
kinetic-energy sum is re-ordered (1e-13)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

from tests.cases import STEP_CASES, build_case, inputs  # noqa: E402
from tests.test_fv_gpu import make_solver  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("name,world", [("journal2d_periodic", 2), ("journal2d_periodic", 3), ("slider2d_xDN", 2),
                                        ("slider2d_allDN_slip", 3), ("journal2d_flip_mc0", 4),
                                        ("journal2d_mcm1_bulk", 2), ("asym_x_mc1", 2), ("asym_y_mcm1_slip", 3),
                                        ("asym_xy_mc0", 2), ("slider2d_allDN_mcm1", 3), ("slider2d_xDN_mc0", 2)])
def test_in_process_slabs_match_single_domain(name, world):
    from gapflow_b200.dist import InProcessSlabs
    d, extra = inputs(name)
    o = build_case(name)
    single, _ = make_solver(name, oracle=o)
    slabs = InProcessSlabs(d, o.q, o.topo, world, extra_global=extra)
    assert np.array_equal(slabs.gather_q(), o.q)
    sc0 = slabs.scalars()
    assert all(s.dt == single.scalars().dt for s in sc0)
    for k in range(5):
        o.update()
        single.step(1)
        slabs.step(1)
        q = slabs.gather_q()
        assert np.array_equal(q, single.get_q()), (name, world, k, int((q != single.get_q()).sum()))
        assert np.array_equal(q, o.q)
        ref = single.scalars()
        for s in slabs.scalars():
            assert s.dt == ref.dt == o.dt and s.step == ref.step
            assert s.v_max == ref.v_max and s.v_sound == ref.v_sound
            assert abs(s.ekin - ref.ekin) <= 1e-13 * abs(ref.ekin)
            assert abs(s.residual - ref.residual) <= 1e-9 * abs(ref.residual) + 1e-14


@pytest.mark.parametrize("world", [1, 2])
@pytest.mark.parametrize("name", ['journal2d_periodic', 'slider2d_xDN', 'slider2d_allDN_slip'])
def test_edge_plus_interior_launches_equal_one_pass(name, world):
    """gpf_fv_pass_part: EDGE (first and last row of tiles) + INTERIOR launches == the single
    launch, bit for bit, on a grid of 11 x 3 ragged tiles (83 x 150), also across 2 slabs."""
    import copy
    from gapflow_b200.dist import InProcessSlabs
    from gapflow_b200.fv import FVSolver
    from oracle.fv import FVOracle
    from tests.cases import STEP_CASES, perturb, slip_field
    d, _ = inputs(name)
    d = copy.deepcopy(d)
    g = d['grid']
    g['Nx'], g['Ny'] = 83, 150
    if name == 'slider2d_xDN':
        g['dx'], g['dy'] = g['Lx'] / 83, g['Ly'] / 150
    else:
        g['Lx'], g['Ly'] = g['dx'] * 83, g['dy'] * 150
    extra = slip_field((85, 152), 1e-6) if STEP_CASES[name].get('extra') else None
    o = FVOracle(g, d['geometry'], d['properties'], d['numerics'], extra=extra)
    perturb(o.q, d['properties']['rho0'], max(abs(d['geometry']['U']), abs(d['geometry']['V'])), 99)
    o.fill_ghosts()
    single = FVSolver(g, d['geometry'], d['properties'], d['numerics'], topo=o.topo, extra=extra)
    single.set_q(o.q)
    single.pre_run(False)
    slabs = InProcessSlabs(d, o.q, o.topo, world, extra_global=extra, split=True)
    for k in range(4):
        single.step(1)
        slabs.step(1)
        assert np.array_equal(slabs.gather_q(), single.get_q()), (name, world, k)
        ref = single.scalars()
        for sc in slabs.scalars():
            assert sc.dt == ref.dt and sc.v_max == ref.v_max and sc.v_sound == ref.v_sound
            assert abs(sc.ekin - ref.ekin) <= 1e-13 * abs(ref.ekin)


@pytest.mark.parametrize("name,world", [("journal2d_periodic", 2), ("journal2d_periodic", 3), ("slider2d_xDN", 2),
                                        ("slider2d_allDN_slip", 3), ("journal2d_flip_mc0", 4), ("journal2d_mcm1_bulk", 2),
                                        ("asym_xy_mc0", 3), ("slider2d_allDN_mcm1", 2), ("journal1d_periodic", 5)])
def test_fused_exchange_matches_single_domain(name, world):
    """gpf_fv_slab_pass: edge rows stored into the neighbours' ghost rows by the pass kernels' boundary tiles, flags and
    record fold by the general launch -- all ranks in one process, driven pass by pass in lock step.  q and dt bit for bit
    equal to the single-domain solver and the oracle (ring seams, physical x sides, every MacCormack order, ragged slabs)."""
    from gapflow_b200.dist import InProcessSlabs
    d, extra = inputs(name)
    o = build_case(name)
    single, _ = make_solver(name, oracle=o)
    slabs = InProcessSlabs(d, o.q, o.topo, world, extra_global=extra, fused=True)
    assert all(s.dt == single.scalars().dt for s in slabs.scalars())
    for k in range(6):
        o.update()
        single.step(1)
        slabs.step(1)
        q, q1 = slabs.gather_q(), single.get_q()
        assert np.array_equal(q, q1), (name, world, k, int((q != q1).sum()))
        if 'EOS: DH' in STEP_CASES[name]['yaml']:
            assert np.array_equal(q, o.q)
        ref = single.scalars()
        for s in slabs.scalars():
            assert s.dt == ref.dt and s.step == ref.step == k + 1
            assert s.v_max == ref.v_max and s.v_sound == ref.v_sound
            assert abs(s.ekin - ref.ekin) <= 1e-13 * abs(ref.ekin)
            assert abs(s.residual - ref.residual) <= 1e-9 * abs(ref.residual) + 1e-14
    slabs.close()


@pytest.mark.parametrize("nx,world,aligned", [(48, 2, True), (48, 3, True), (37, 2, False), (50, 3, False)])
def test_fused_exchange_fast_contract(nx, world, aligned):
    """The contracted arithmetic is one kernel binary, but a cell's update is compiled in several instances (first /
    second row of a thread, own / halo cell of a tile) whose FMA contraction may round the last bit differently.  With
    slab heights that are multiples of the 8-row tile every cell plays the same role as in the single domain and the
    fields are bit-identical; otherwise they agree to the contract's tolerance (1e-12 per step)."""
    import copy
    from gapflow_b200.dist import InProcessSlabs
    from gapflow_b200.fv import FVSolver
    from oracle.fv import FVOracle
    from tests.cases import perturb
    from tests.test_fv_gpu import relerr
    d, _ = inputs('journal2d_periodic')
    d = copy.deepcopy(d)
    g = d['grid']
    g['Nx'], g['Lx'] = nx, g['dx'] * nx
    o = FVOracle(g, d['geometry'], d['properties'], d['numerics'])
    perturb(o.q, d['properties']['rho0'], 0.1, 5)
    o.fill_ghosts()
    single = FVSolver(g, d['geometry'], d['properties'], d['numerics'], topo=o.topo, arith='fast')
    single.set_q(o.q)
    single.pre_run(False)
    slabs = InProcessSlabs(d, o.q, o.topo, world, fused=True, arith='fast')
    for k in range(5):
        single.step(1)
        slabs.step(1)
        q, q1 = slabs.gather_q(), single.get_q()
        if aligned:
            assert np.array_equal(q, q1), (nx, world, k, int((q != q1).sum()))
            assert slabs.scalars()[0].dt == single.scalars().dt
        else:
            assert relerr(q, q1) <= 1e-12 * (k + 1)
    slabs.close()


def test_fused_exchange_invalid_step_and_general_kernel_redo():
    """a step that goes invalid stops every rank with the previous field in place; operands outside the straight-line
    range make the general launch recompute the pass -- it then re-sends the edge rows and closes the pass itself"""
    from gapflow_b200.dist import InProcessSlabs
    name = "journal2d_periodic"
    d, extra = inputs(name)
    o = build_case(name)
    slabs = InProcessSlabs(d, o.q, o.topo, 3, fused=True)
    slabs.step(1)
    q1 = slabs.gather_q()
    for s in slabs.s:
        s.fv.set_scalars(dt=1e-3)
    slabs.step(1)
    for s in slabs.s:
        sc = s.fv.scalars()
        assert sc.stop == 1 and sc.step == 1
    assert np.array_equal(slabs.gather_q(), q1)
    slabs.step(2)                                   # halted: the exchanges and folds still run, nothing moves
    assert np.array_equal(slabs.gather_q(), q1)
    slabs.close()
    # tiny momenta (outside [2^-100, 2^100]) in one slab: the straight-line launch flags the pass, the general one redoes it
    o = build_case(name)
    o.q[1, 5:9, 3:7] = 1e-40
    single, _ = make_solver(name, oracle=o)
    slabs = InProcessSlabs(d, o.q, o.topo, 2, fused=True)
    for k in range(3):
        single.step(1)
        slabs.step(1)
        assert np.array_equal(slabs.gather_q(), single.get_q()), k
    slabs.close()


def test_in_process_slabs_invalid_step_stops_every_rank():
    from gapflow_b200.dist import InProcessSlabs
    name = "journal2d_periodic"
    d, extra = inputs(name)
    o = build_case(name)
    slabs = InProcessSlabs(d, o.q, o.topo, 2)
    slabs.step(1)
    q1 = slabs.gather_q()
    for s in slabs.s:
        s.fv.set_scalars(dt=1e-3)
    slabs.step(1)
    for s in slabs.s:
        sc = s.fv.scalars()
        assert sc.stop == 1 and sc.step == 1
    assert np.array_equal(slabs.gather_q(), q1)


NCCL_WORKER = r"""
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
from tests.cases import build_case, inputs
from gapflow_b200.dist import SlabSolver
for name in ("journal2d_periodic", "slider2d_xDN", "slider2d_allDN_slip"):
    d, extra = inputs(name)
    o = build_case(name)
    s = SlabSolver.from_global(d, o.q, o.topo, extra)
    for k in range(6):
        o.update()
    s.step(6)
    q = s.gather_q()
    sc = s.scalars()
    assert np.array_equal(q, o.q), (name, rank, int((q != o.q).sum()))
    assert sc.dt == o.dt and sc.step == o.step, (name, sc.dt, o.dt)
    assert abs(sc.ekin - o.kinetic_energy) <= 1e-13 * abs(o.kinetic_energy)
    m = s.mass
    assert abs(m - o.mass) <= 1e-12 * abs(o.mass)
# GP inference with the query cells partitioned over the ranks (training set replicated)
from gapflow_b200.gp import GPModel, features
from gapflow_b200.dist import PartitionedGP, owned_rows, slab_rows
from oracle import gp as G
from tests.gp_cases import training_set
name = "slider2d_allDN_slip"
d, extra = inputs(name)
o = build_case(name)
s = SlabSolver.from_global(d, o.q, o.topo, extra)
c = training_set(150, 'press_wide')
xs = np.array([o.q[0].max() / 0.9, 1., 1., 1., 1., 1., 1.])      # maps rho into the training range
gp = PartitionedGP(GPModel(200, 1, 1))
gp.factor(c['X'], c['Y'], c['log_amp'], c['log_scale'], c['Yerr'])
Xloc = features(s.fv, xs, [0])
Ny2 = d['grid']['Ny'] + 2
a, b = owned_rows(d['grid']['Nx'], rank, world)
start = slab_rows(d['grid']['Nx'], world)[rank][0]
Xown = Xloc[(a - start) * Ny2:(b - start) * Ny2].contiguous()
m, v = gp.mean_var(Xown, a * Ny2, 1.5e-8)
Xg = G.features(o.q, o.topo, o.extra, xs, [0])
L = G.build(c['X'], c['Yerr'], c['log_amp'], c['log_scale'])
al = G.alpha(L, c['Y'])
m_ref, v_ref = G.repredict_mean_var(c['X'], c['log_amp'], c['log_scale'], L, al, 1.5e-8, Xg)
assert np.abs(v.cpu().numpy() - v_ref[a * Ny2:b * Ny2]).max() <= 1e-9
assert np.abs(m.cpu().numpy()[0] - m_ref[0, a * Ny2:b * Ny2]).max() <= 1e-9 * np.abs(m_ref).max()
assert gp.argmax == int(np.argmax(v_ref)), (gp.argmax, int(np.argmax(v_ref)))
assert abs(gp.maximum_variance - v_ref.max()) <= 1e-9
dist.barrier()
if rank == 0:
    print("NCCL_SLABS_OK")
dist.destroy_process_group()
"""


def test_nccl_slabs_match_oracle(tmp_path):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(n, 4)
    r = _run_workers(tmp_path, NCCL_WORKER, world)
    assert r.returncode == 0 and "NCCL_SLABS_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


# Real processes, one GPU each, handles over gloo instead of NCCL: the CUDA-IPC mapping, the system-scope
# release / acquire flags, the slot parity and the graph replay of gpf_fv_slab_step with no NCCL anywhere,
# plus the lagging-peer report.  Needs one GPU per rank: kernels that wait on one another must never share a
# device (two such processes on one B200 end in a context-switch timeout, Xid 109), so on the 1-GPU test box
# these skip and the multi-process path is proven by bench.py's `slab_parity` block in every N > 1 run.
IPC_WORKER = r"""
import os, sys, time, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
mode = sys.argv[2]
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
assert torch.cuda.device_count() >= world, "one GPU per rank"
torch.cuda.set_device(rank)
dist.init_process_group("gloo")
from tests.cases import build_case, inputs
from gapflow_b200 import _capi
from gapflow_b200.dist import SlabSolver
if mode == "parity":
    for name, arith in (("journal2d_periodic", "exact"), ("slider2d_xDN", "exact"), ("slider2d_allDN_slip", "exact"),
                        ("journal2d_mcm1_bulk", "exact"), ("journal2d_periodic", "fast")):
        d, extra = inputs(name)
        o = build_case(name)
        s = SlabSolver.from_global(d, o.q, o.topo, extra, peer=True, arith=arith)
        assert s._comm is not None
        for k in range(7):
            o.update()
        s.step(3)                       # first call: plain launches + graph capture
        s.step(4)                       # replays
        q = s.gather_q()
        sc = s.scalars()
        if arith == "exact":
            assert np.array_equal(q, o.q), (name, rank, int((q != o.q).sum()))
            assert sc.dt == o.dt and sc.step == o.step, (name, sc.dt, o.dt)
        else:
            assert np.abs(q - o.q).max() <= 1e-11 * np.abs(o.q).max(), name
        assert abs(sc.ekin - o.kinetic_energy) <= 1e-11 * abs(o.kinetic_energy)
        assert abs(s.mass - o.mass) <= 1e-12 * abs(o.mass)
        assert not s.comm_error()
        s.close()
    dist.barrier()
    if rank == 0:
        print("IPC_SLABS_OK")
else:
    # a lagging peer: rank 1 does not step; rank 0's waits exceed the bound, the run stops and the host is told
    name = "journal2d_periodic"
    d, extra = inputs(name)
    o = build_case(name)
    s = SlabSolver.from_global(d, o.q, o.topo, extra, peer=True)
    s.set_comm_timeout(1.0)
    dist.barrier()
    if rank == 0:
        s.step(2)
        try:
            s.scalars()
            print("NO_ERROR_RAISED")
        except _capi.GapflowError as e:
            assert e.code == _capi.GPF_E_STATE, e
            assert s.fv.lib.gpf_last_error()
            print("LAG_REPORTED")
    dist.barrier()
    torch.cuda.synchronize()
sys.stdout.flush()
os._exit(0)
"""


def _run_workers(tmp_path, source, world, *args, timeout=600):
    script = tmp_path / "worker.py"
    script.write_text(source)
    sock = socket.socket()
    sock.bind(("127.0.0.1", 0))
    port = sock.getsockname()[1]
    sock.close()
    return subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                           "--master-addr", "127.0.0.1", "--master-port", str(port), str(script), ROOT, *args],
                          capture_output=True, text=True, timeout=timeout, cwd=ROOT)


@pytest.mark.parametrize("world", [2, 3])
def test_ipc_slabs_in_real_processes_match_oracle(tmp_path, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs >= {world} GPUs (one per rank)")
    r = _run_workers(tmp_path, IPC_WORKER, world, "parity")
    assert r.returncode == 0 and "IPC_SLABS_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_ipc_lagging_peer_is_reported_not_silent(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs (one per rank)")
    r = _run_workers(tmp_path, IPC_WORKER, 2, "lag")
    assert r.returncode == 0 and "LAG_REPORTED" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
