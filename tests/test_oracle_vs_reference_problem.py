"""CPU, where the reference tree is present: oracle/fv.py against the reference's own ``Problem`` class, EXECUTED
(oracle/ref_problem.py imports GaPFlow/problem.py, models/stress.py, topography.py, integrate.py ... unchanged, with inert
stand-ins for the packages that are not installed; none of them computes anything for fixed-form closures).

This is the pin of the update LOOP itself -- ``update``, the MacCormack order switch, ``_communicate_ghost_buffers`` /
``_get_ghost_cell_values`` with their mask cross-lookups, the stress models' scatter and halving, the viscosity
modifiers, ``_post_update`` (residual, dt), ``_finalize`` -- which tests/test_oracle_vs_reference.py pins only through
the imported flux / closure functions.  Every golden case of tests/cases.py, bit for bit."""
import contextlib
import io

import numpy as np
import pytest

from oracle import ref_loader
from tests import cases

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")
STEPS = 6


@pytest.fixture(scope="module")
def RefProblem():
    from oracle import ref_problem
    yield ref_problem.reference_problem()
    ref_problem.release()


def _reference(RefProblem, name, q_start=None):
    """The reference's Problem for a case of tests/cases.py, its own YAML reader, `_pre_run` done."""
    import sys
    read_yaml_input = sys.modules['GaPFlow.io'].read_yaml_input
    _, extra = cases.inputs(name)
    with io.StringIO(cases.STEP_CASES[name]['yaml']) as f, contextlib.redirect_stdout(io.StringIO()):
        d = read_yaml_input(f)
        d['options']['silent'] = True
        p = RefProblem(*RefProblem._get_mandatory_input(d), extra_field=extra)
        p._pre_run()
    if q_start is not None:
        p.q[...] = q_start
        if d['numerics']['adaptive']:
            p.dt = d['numerics']['CFL'] * p.dt_crit          # what build_case does on the oracle
    return p


def _same_state(o, p, name, k):
    assert np.array_equal(o.q, p.q), (name, k)
    assert o.dt == p.dt and o.step == p.step and o.simtime == p.simtime and o.residual == p.residual, (name, k)
    assert o.mass == p.mass and o.kinetic_energy == p.kinetic_energy and o.v_max == p.v_max, (name, k)
    assert o.v_sound == p.pressure.v_sound and o.dt_crit == p.dt_crit and o.cfl == p.cfl, (name, k)
    assert list(o.residual_buffer) == list(p.residual_buffer) and o.converged == bool(p.converged), (name, k)
    # closure fields as they stand after a step: those of the corrector pass's input (problem.py:532-540 vs :563)
    assert np.array_equal(o.pressure, p.pressure.pressure), (name, k)
    assert np.array_equal(o.wall_xz, p.wall_stress_xz.full) and np.array_equal(o.wall_yz, p.wall_stress_yz.full), (name, k)
    assert np.array_equal(o.bulk, p.bulk_stress.stress), (name, k)


@pytest.mark.parametrize('name', list(cases.STEP_CASES))
def test_update_loop_bit_for_bit(RefProblem, name):
    o = cases.build_case(name)
    p = _reference(RefProblem, name, q_start=o.q)
    assert np.array_equal(o.topo[:3], p.topo.full[:3])
    assert np.array_equal(o.extra, p.pressure._GaussianProcessSurrogate__extra.p)
    assert o.dt == p.dt and o.kinetic_energy_old == p.kinetic_energy_old
    for k in range(STEPS):
        o.update()
        p.update()
        _same_state(o, p, name, k)


@pytest.mark.parametrize('name', list(cases.STEP_CASES))
def test_ghost_cells_bit_for_bit(RefProblem, name):
    """fill on a field whose ghost cells hold garbage: every ghost value must come from the rule, none from the input"""
    o = cases.build_case(name)
    p = _reference(RefProblem, name)
    rng = np.random.default_rng(5)
    q = o.q.copy()
    for sl in ((slice(None), 0), (slice(None), -1), (slice(None), slice(None), 0), (slice(None), slice(None), -1)):
        q[sl] = rng.uniform(-1e3, 1e3, q[sl].shape)
    o.q[...] = q
    p.q[...] = q
    o.fill_ghosts()
    p._communicate_ghost_buffers()
    assert np.array_equal(o.q, p.q), name


def test_invalid_state_takes_the_reference_s_exit(RefProblem):
    """_finalize (problem.py:588-610): a negative density restores the previous field, recomputes the closures, stops"""
    name = 'slider2d_xDN'
    o = cases.build_case(name)
    p = _reference(RefProblem, name, q_start=o.q)
    for x in (o, p):
        x.update()
        x.dt = 1e-3                   # far beyond the stability limit
    with contextlib.redirect_stdout(io.StringIO()):
        for _ in range(3):
            o.update()
            p.update()
            if o._stop:
                break
    assert o._stop and p._stop
    assert np.array_equal(o.q, p.q) and o.step == p.step
    assert np.array_equal(o.pressure, p.pressure.pressure) and np.array_equal(o.wall_xz, p.wall_stress_xz.full)


@pytest.mark.parametrize('name', list(cases.STEP_CASES))
def test_committed_golden_vectors_are_what_the_reference_computes(RefProblem, golden_dir, name):
    """tests/golden/fv_steps.npz (what the GPU parity tests compare with on the box, where the reference tree is absent)
    against the executed reference Problem: start field, dt, fields and scalars after 1 and 3 steps."""
    import os
    gs = np.load(os.path.join(golden_dir, "fv_steps.npz"))
    bwr = 'EOS: BWR' in cases.STEP_CASES[name]['yaml']          # np.power / np.exp: CPU-dispatched SIMD routines, ~1 ulp
    same = (lambda a, b: np.abs(a - b).max() <= 1e-13 * np.abs(b).max()) if bwr else np.array_equal
    p = _reference(RefProblem, name, q_start=gs[f'{name}_q0'])
    assert p.dt == gs[f'{name}_dt0'] or bwr
    for n, tag in ((1, '1'), (2, '3')):
        for _ in range(n):
            p.update()
        assert same(p.q, gs[f'{name}_q{tag}']), (name, tag)
        s = np.array([p.dt, p.residual, p.kinetic_energy, p.v_max, p.pressure.v_sound, p.mass])
        assert same(s, gs[f'{name}_s{tag}']), (name, tag)


@pytest.mark.parametrize('tol, max_it', [(1e-4, 10_000), (1e-12, 150)])
def test_run_loop_stops_where_the_reference_s_does(RefProblem, tol, max_it):
    """Problem.run (problem.py:368-410): until the last five residuals are below `tol`, or `max_it` steps"""
    import sys
    from gapflow_b200 import config
    from oracle.fv import FVOracle
    from tests.test_problem_gpu import SOMMERFELD
    text = SOMMERFELD.replace("tol: 1e-8", f"tol: {tol}").replace("max_it: 10_000", f"max_it: {max_it}")
    d = config.read_yaml_string(text)
    o = FVOracle(d['grid'], d['geometry'], d['properties'], d['numerics'])
    o.run()
    with io.StringIO(text) as f, contextlib.redirect_stdout(io.StringIO()):
        rd = sys.modules['GaPFlow.io'].read_yaml_input(f)
        rd['options']['silent'] = True
        p = RefProblem._from_dict(rd)
        p.run()
    assert o.step == p.step and (o.step < max_it) == (tol > 1e-6) and o.step > 50
    assert np.array_equal(o.q, p.q) and o.dt == p.dt and o.simtime == p.simtime
    assert list(o.residual_buffer) == list(p.residual_buffer)
