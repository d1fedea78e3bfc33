"""CPU, travels: the GP-coupled step's oracle (oracle/surrogate.py) against tests/golden/gp_coupled.npz, which
oracle/make_golden_gp.py wrote from the EXECUTED reference (its own Problem / surrogate / database / mock-MD classes;
tinygp replaced by the stated stand-in, so GP parity stays unpinned at that boundary).  The GPU parity tests compare the
CUDA path with this oracle on the box, where the reference tree is absent."""
import contextlib
import io
import os

import numpy as np
import pytest

from oracle import surrogate as OS
from oracle.make_golden_gp import CASES, FIELD
from tests import surrogate_cases as SC


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "gp_coupled.npz"))


def _rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-300)


def _check(o, odb, G, tag):
    dt, residual, v_sound, size, step = G(f's_{tag}')
    assert odb.size == size and o.step == step, tag
    assert _rel(o.q, G(f'q_{tag}')) <= 1e-12 and abs(o.dt - dt) <= 1e-12 * dt, tag
    assert abs(o.residual - residual) <= 1e-9 * max(abs(residual), 1e-300), tag
    assert abs(o.v_sound - v_sound) <= 1e-9 * v_sound, tag
    assert _rel(o.pressure, G(f'p_{tag}')) <= 1e-12, tag
    assert _rel(o.wall_xz, G(f'wxz_{tag}')) <= 1e-12 and _rel(o.wall_yz, G(f'wyz_{tag}')) <= 1e-12, tag
    for k, m in o.models.items():
        vmax, vtol, pause, nfit, amp, yscale, yerr = G(f'm_{k}_{tag}')
        unit = amp * yscale**2
        assert np.abs(o.variance[k] - G(f'var_{k}_{tag}')).max() <= 1e-9 * unit, (tag, k)
        assert abs(m.maximum_variance - vmax) <= 1e-9 * unit and abs(m.variance_tol - vtol) <= 1e-12 * vtol, (tag, k)
        assert m._pause == pause and m.last_fit_train_size == nfit, (tag, k)
        assert abs(np.exp(m.log_amp) - amp) <= 1e-14 and abs(m.Yscale - yscale) <= 1e-14 * yscale, (tag, k)
        assert abs(m.Yerr - yerr) <= 1e-14 * yerr, (tag, k)


@pytest.mark.parametrize('case', list(CASES))
def test_oracle_reproduces_the_executed_reference(gold, case):
    name, kw, steps, al = CASES[case]
    G = lambda key: gold[f'{case}/{key}']           # noqa: E731
    d, _ = SC.inputs(name, **kw)
    if not al:
        for k in ('press', 'shear'):
            d['gp'][k]['active_learning'] = False
    g = d['grid']
    odb = OS.Database(OS.MockMD(d['properties'], d['geometry'], d['gp'], draws=SC.DRAWS), d['db'])
    o = OS.SurrogateFVOracle(g, d['geometry'], d['properties'], d['numerics'], d['gp'], odb,
                             extra=SC.slip_length(g['Nx'], g['Ny']), write_freq=d['options']['write_freq'],
                             optimizer=SC.fixed_optimizer)
    o.q[...] = SC.perturb(o.q)
    o.fill_ghosts()
    o.pre_run()
    _check(o, odb, G, 'pre')
    for k in range(steps):
        with contextlib.redirect_stdout(io.StringIO()):
            o.update()
        _check(o, odb, G, str(k))
    assert _rel(odb._Xtrain, G('X')) <= 1e-12 and _rel(odb._Xtrain[:10], G('X')[:10]) <= 1e-14
    assert _rel(odb._Ytrain, G('Y')) <= 1e-13 and np.array_equal(odb._Ytrain_err, G('Yerr'))
    assert (odb.size > 10) == al
    for k, m in o.models.items():
        assert FIELD[k]
        for key, h in m.history.items():
            want = G(f'hist_{k}_{key}')
            assert len(h) == len(want) and (len(h) == 0 or _rel(np.array(h, dtype=float), want) <= 1e-9), (k, key)
