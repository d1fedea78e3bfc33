"""CPU, world_size 2, 3, 4 and 8 over gloo: the slab partition, neighbour table, message pairing and
rank-ordered scalar fold of gapflow_b200/dist.py reproduce the single-domain result
bit-for-bit (q, dt) when they drive the NumPy oracle on slabs (tests/slab_oracle.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gapflow_b200.dist import HaloExchanger, slab_rows
from tests.cases import build_case, inputs

NSTEPS = 4


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tests.slab_oracle import SlabOracle
        d, extra = inputs(name)
        ref = build_case(name)                     # every rank builds the same global start state
        s = SlabOracle(d, ref.q, ref.topo, extra, rank, world)
        dts = [s.dt]
        for _ in range(NSTEPS):
            s.update()
            dts.append(s.dt)
        q = s.gather_q(d['grid']['Nx'])
        if rank == 0:
            np.savez(out_path, q=q, dts=np.array(dts), ekin=s.kinetic_energy_old, step=s.step, res=s.residual)
        else:
            _ = s.kinetic_energy_old
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name,world", [("journal2d_periodic", 2), ("slider2d_xDN", 2), ("slider2d_allDN_slip", 3),
                                        ("journal2d_flip_mc0", 2),
                                        # the rank counts of the scaling run: a ring of 4 and of 8 (uneven slabs: 37 rows), and
                                        # different masks on the two x sides held by different ranks
                                        ("journal2d_periodic", 4), ("journal2d_periodic", 8), ("asym_x_mc1", 4)])
def test_slab_oracle_matches_single_domain(tmp_path, name, world):
    out = str(tmp_path / "out.npz")
    mp.spawn(_worker, args=(world, _free_port(), name, out), nprocs=world, join=True)
    got = np.load(out)
    ref = build_case(name)
    dts = [ref.dt]
    for _ in range(NSTEPS):
        ref.update()
        dts.append(ref.dt)
    assert int(got['step']) == NSTEPS
    assert np.array_equal(got['q'], ref.q), (name, np.abs(got['q'] - ref.q).max())
    assert np.array_equal(got['dts'], np.array(dts))            # maxima are order-independent
    assert abs(float(got['ekin']) - ref.kinetic_energy_old) <= 1e-13 * abs(ref.kinetic_energy_old)
    assert abs(float(got['res']) - ref.residual) <= 1e-9 * abs(ref.residual) + 1e-14


def test_slab_rows_cover_the_grid():
    for Nx in (1, 7, 64, 100, 4097):
        for w in (1, 2, 3, 8):
            rows = slab_rows(Nx, w)
            assert sum(n for _, n in rows) == Nx
            assert rows[0][0] == 0 and all(rows[i][0] + rows[i][1] == rows[i + 1][0] for i in range(w - 1))
            assert max(n for _, n in rows) - min(n for _, n in rows) <= 1


def test_neighbour_table():
    assert (HaloExchanger(0, 1, True).lo, HaloExchanger(0, 1, True).hi) == (None, None)
    assert (HaloExchanger(0, 4, True).lo, HaloExchanger(0, 4, True).hi) == (3, 1)
    assert (HaloExchanger(3, 4, True).lo, HaloExchanger(3, 4, True).hi) == (2, 0)
    assert (HaloExchanger(0, 4, False).lo, HaloExchanger(3, 4, False).hi) == (None, None)
    assert (HaloExchanger(1, 2, True).lo, HaloExchanger(1, 2, True).hi) == (0, 0)


def _argmax_worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from gapflow_b200.dist import fold_argmax, owned_rows
        Nx, Ny2 = 10, 7
        rng = np.random.default_rng(5)
        var = rng.random((Nx + 2) * Ny2)
        var[[11, 40, 70]] = 2.0                       # the maximum occurs three times: the first index wins
        a, b = owned_rows(Nx, rank, world)
        loc = var[a * Ny2:b * Ny2]
        v, i = fold_argmax(torch.tensor(loc.max()), torch.tensor(float(np.argmax(loc) + a * Ny2)))
        vn = var.copy()
        vn[55] = np.nan
        locn = vn[a * Ny2:b * Ny2]
        has = np.isnan(locn).any()
        vv = torch.tensor(float('nan') if has else locn.max())
        ii = torch.tensor(float((np.argmax(np.isnan(locn)) if has else np.argmax(locn)) + a * Ny2))
        v2, i2 = fold_argmax(vv, ii)
        if rank == 0:
            np.savez(out_path, v=v, i=i, i2=i2, nan=np.isnan(v2))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_partitioned_argmax_is_the_global_first_maximum(tmp_path, world):
    out = str(tmp_path / "am.npz")
    mp.spawn(_argmax_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    got = np.load(out)
    assert float(got['v']) == 2.0 and int(got['i']) == 11
    assert bool(got['nan']) and int(got['i2']) == 55


def test_owned_rows_partition_every_row_once():
    from gapflow_b200.dist import owned_rows
    for Nx in (5, 64, 101):
        for w in (1, 2, 3, 8):
            if w > Nx:
                continue
            rows = [owned_rows(Nx, r, w) for r in range(w)]
            assert rows[0][0] == 0 and rows[-1][1] == Nx + 2
            assert all(rows[r][1] == rows[r + 1][0] for r in range(w - 1))


# ---- host-level folds of the slab-decomposed GP-coupled step (gapflow_b200/surrogate.py) -------------------
def test_fold_first_maximum_matches_numpy_argmax():
    """(max, argmax) of a global array from per-rank (local max, global index of its first occurrence), with
    overlapping (halo) rows, ties and NaN -- jnp.max / np.argmax semantics (gp.py:406, 424)."""
    from gapflow_b200.surrogate import fold_first_maximum
    rng = np.random.default_rng(5)
    for trial in range(50):
        n, w = 40, 4
        v = np.round(rng.random(n), 1)                       # many ties
        if trial % 7 == 0:
            v[rng.integers(n)] = np.nan
        bounds = [(max(0, r * 10 - 1), min(n, (r + 1) * 10 + 1)) for r in range(w)]     # one overlapping cell per side
        pairs = []
        for a, b in bounds:
            loc = v[a:b]
            k = int(np.argmax(loc))                          # np.argmax: first maximum, NaN wins
            pairs.append((float(loc[k]), a + k))
        vmax, imax = fold_first_maximum(pairs)
        assert imax == int(np.argmax(v)), (trial, pairs)
        assert (np.isnan(vmax) and np.isnan(np.max(v))) or vmax == np.max(v)


def test_thread_collective_allgather_and_exchanger():
    """ThreadColl: rank-ordered all-gather between threads and the halo / record exchange built on it (CPU tensors)."""
    import threading
    import torch
    from gapflow_b200.surrogate import ThreadColl
    world = 3
    tc = ThreadColl(world)
    out = [None] * world

    def work(r):
        c = tc.view(r)
        got = [c.allgather((r, k)) for k in range(5)]
        ex = tc.exchanger(r, periodic=True)
        send = torch.full((2, 3, 4), float(r), dtype=torch.float64)
        send[1] += 0.5                                       # [0] goes down, [1] goes up
        recv = torch.zeros_like(send)
        ex.exchange(send, recv)
        recs = torch.zeros((world, 5), dtype=torch.float64)
        ex.gather(torch.full((5,), float(r), dtype=torch.float64), recs)
        out[r] = (got, recv, recs)

    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join(60) for t in th]
    for r in range(world):
        got, recv, recs = out[r]
        assert got == [[(q, k) for q in range(world)] for k in range(5)]
        lo, hi = (r - 1) % world, (r + 1) % world
        assert torch.all(recv[0] == lo + 0.5) and torch.all(recv[1] == float(hi))      # lower neighbour's last row, upper's first
        assert torch.equal(recs, torch.arange(world, dtype=torch.float64)[:, None].expand(world, 5))


def _gloo_coll_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from gapflow_b200.surrogate import DistColl, fold_first_maximum
    c = DistColl()
    pairs = c.allgather((float(rank % 2), 10 * rank))
    rows = c.allgather(None if rank == 0 else np.full(7, float(rank)))
    q.put((rank, fold_first_maximum(pairs), next(r for r in rows if r is not None)[0]))
    dist.destroy_process_group()


def test_dist_collective_over_gloo():
    """DistColl (torch.distributed all_gather_object) gives every rank the same fold: world_size 2, gloo."""
    import socket
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_gloo_coll_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=120) for _ in ps)
    [p.join(60) for p in ps]
    assert res == [(0, (1.0, 10), 1.0), (1, (1.0, 10), 1.0)]
