"""GPU: members of the reference's ``Problem`` that scripts reach for besides update / run (problem.py:319-331 validity
flags over the padded array, :676-707 ``_communicate_ghost_buffers`` after writing into ``q``).  Collected last."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

from gapflow_b200 import config  # noqa: E402
from tests import cases  # noqa: E402


def test_validity_flags_and_ghost_fill_after_a_host_edit():
    from gapflow_b200.problem import Problem
    name = 'slider2d_xDN'                     # Dirichlet / Neumann in x, periodic in y: every ghost rule, corners
    o = cases.build_case(name)                # oracle: perturbed interior, ghost cells filled, dt set
    d = config.read_yaml_string(cases.STEP_CASES[name]['yaml'])
    p = Problem._from_dict(d)
    assert p.q_is_valid and not p.q_has_nan and not p.q_has_negative_density
    rng = np.random.default_rng(5)
    q = o.q.copy()
    for sl in ((slice(None), 0), (slice(None), -1), (slice(None), slice(None), 0), (slice(None), slice(None), -1)):
        q[sl] = rng.uniform(1., 2., q[sl].shape)           # garbage in the ghost cells
    p.q[...] = q
    p._communicate_ghost_buffers()
    assert np.array_equal(p.q, o.q)
    p._pre_run()
    assert p.dt == o.dt
    for _ in range(2):
        p.update()
        o.update()
    assert np.array_equal(p.q, o.q) and p.q_is_valid
    p.q[0, 3, 4] = -1.
    assert p.q_has_negative_density and not p.q_is_valid and not p.q_has_nan
    p.q[0, 3, 4] = np.nan
    assert p.q_has_nan and not p.q_is_valid
