"""Inputs of the GP-coupled step tests: the reference's mock-MD examples (examples/slip_1d_lj_mock.py,
examples/slip_2d_lj_mock.py) scaled down, in the reference's YAML schema, plus the explicit inputs that
replace what the reference draws from JAX (see oracle/surrogate.py)."""
import numpy as np

from gapflow_b200 import config

SLIP_1D = """
options: {{output: {out}, write_freq: {wf}, use_tstamp: False, silent: {silent}}}
grid: {{Lx: 1470., Ly: 1., Nx: {nx}, Ny: 1, xE: ['P','P','P'], xW: ['P','P','P'], yS: ['P','P','P'], yN: ['P','P','P']}}
geometry: {{type: inclined, hmin: 12., hmax: 12., U: 0.12, V: 0.}}
numerics: {{CFL: 0.5, adaptive: 1, tol: 1e-8, dt: 0.1, max_it: {max_it}}}
properties: {{shear: 2.15, bulk: 0., EOS: BWR, T: 1.0, rho0: 0.8}}
gp:
  press: {{fix_noise: True, atol: {atol}, rtol: 0., obs_stddev: 2.e-2, max_steps: {al_steps}, pause_steps: 3, active_dims: [0]}}
  shear: {{fix_noise: True, atol: {atol}, rtol: 0., obs_stddev: 4.e-3, max_steps: {al_steps}, pause_steps: 3, active_dims: {{x: {sx}}}}}
db: {{init_size: {init}, init_method: lhc}}
"""

# 2D: journal gap, Dowson-Higginson-free (BWR) fluid of the mock example, all three GPs
SLIP_2D = """
options: {{output: {out}, write_freq: {wf}, use_tstamp: False, silent: {silent}}}
grid: {{Lx: 1470., Ly: 1000., Nx: {nx}, Ny: {ny}, xE: ['P','P','P'], xW: ['P','P','P'], yS: ['P','P','P'], yN: ['P','P','P']}}
geometry: {{type: journal, hmin: 10., hmax: 14., U: 0.12, V: 0.02}}
numerics: {{CFL: 0.4, adaptive: 1, tol: 1e-8, dt: 0.1, max_it: {max_it}}}
properties: {{shear: 2.15, bulk: 0.5, EOS: BWR, T: 1.0, rho0: 0.8}}
gp:
  press: {{fix_noise: True, atol: {atol}, rtol: 0., obs_stddev: 2.e-2, max_steps: {al_steps}, pause_steps: 3, active_dims: [0, 3]}}
  shear: {{fix_noise: True, atol: {atol}, rtol: 0., obs_stddev: 4.e-3, max_steps: {al_steps}, pause_steps: 3, active_dims: {{x: [0, 1, 3, 6], y: [0, 2, 3, 6]}}}}
db: {{init_size: {init}, init_method: lhc, init_width: 0.02}}
"""

DRAWS = (0.31, -0.27, 0.54)        # stand-ins for the three jax.random.key(123) normals of md/mock.py:82-88


def inputs(name, nx=100, ny=1, wf=1000, max_it=50, atol=1., al_steps=10, init=10, out='/tmp/gpf_sur', silent=True,
           sx='[0, 1, 6]'):
    text = {'slip1d': SLIP_1D, 'slip2d': SLIP_2D}[name].format(nx=nx, ny=ny, wf=wf, max_it=max_it, atol=atol,
                                                                al_steps=al_steps, init=init, out=out, silent=silent, sx=sx)
    return config.read_yaml_string(text), text


def slip_length(nx, ny):
    """examples/slip_1d_lj_mock.py:79-92: smooth periodic step of the slip length, ghost cells filled."""
    from scipy.special import erf
    sl = np.zeros(nx)
    e = erf(np.linspace(-20., 20., nx // 2))
    sl[:nx // 2] = e
    sl[nx // 2:nx // 2 + len(e)] = -e
    sl = (1. + np.roll(sl, nx // 4)) / 2.
    extra = np.zeros((1, nx + 2, ny + 2))
    extra[0, 1:-1, :] = sl[:, None]
    if ny > 1:                         # a gentle y modulation so that 2D features differ along y
        extra[0, 1:-1, 1:-1] *= 1. + 0.2 * np.cos(2. * np.pi * (np.arange(ny) + 0.5) / ny)[None, :]
        extra[0, :, 0] = extra[0, :, -2]
        extra[0, :, -1] = extra[0, :, 1]
    extra[0, 0, :] = extra[0, -2, :]
    extra[0, -1, :] = extra[0, 1, :]
    return extra


def fixed_optimizer(model):
    """hyperparameters as an explicit function of the training inputs (identical for oracle and product):
    params_init (stress.py:281-284) with the length scales widened by e^0.5."""
    X = np.asarray(model.Xtrain)
    return 0.3, np.log(np.std(X, axis=0)) + 0.5


def perturb(q, seed=7, amp=2e-3):
    """seeded smooth perturbation of the uniform initial state (interior), so that fluxes are non-trivial."""
    rng = np.random.default_rng(seed)
    nx, ny = q.shape[1] - 2, q.shape[2] - 2
    x = (np.arange(nx) + 0.5) / nx
    y = (np.arange(ny) + 0.5) / ny
    out = q.copy()
    for c in range(3):
        a, b = rng.uniform(0.5, 1.0, 2)
        out[c, 1:-1, 1:-1] *= 1. + amp * (a * np.sin(2 * np.pi * x)[:, None] + b * np.cos(2 * np.pi * y)[None, :] * (ny > 1))
    return out
