"""Shared test inputs: YAML texts of the BASELINE configs (scaled) and edge cases.

The YAML strings use the reference's schema verbatim (README.md:73-107,
examples/config/journal_1d_dowson-higginson.yaml, tests/test_flip_axes.py:30-65).
``build_case(name)`` returns a CPU oracle primed with the seeded perturbation of
SURVEY.md section 8(d) config 2, so both flux directions see non-trivial data.
"""
import numpy as np

from gapflow_b200 import config

_DH = """
properties:
    shear: 0.0794
    bulk: {bulk}
    EOS: DH
    P0: 101325.
    rho0: 877.7007
    T0: 323.15
    C1: {C1}
    C2: 1.23
"""


def _yaml(grid, geo, numerics, prop=None, bulk=0., C1='3.5e10'):
    return ("options:\n    output: data/test\n    write_freq: 1000\n    silent: True\n"
            + grid + geo + numerics + (prop if prop is not None else _DH.format(bulk=bulk, C1=C1)))


_PER = "    xE: ['P', 'P', 'P']\n    xW: ['P', 'P', 'P']\n    yS: ['P', 'P', 'P']\n    yN: ['P', 'P', 'P']\n"
_XDN = ("    xE: ['D', 'N', 'N']\n    xW: ['D', 'N', 'N']\n    yS: ['P', 'P', 'P']\n    yN: ['P', 'P', 'P']\n"
        "    xE_D: 877.7007\n    xW_D: 877.7007\n")
_JOURNAL = "geometry:\n    type: journal\n    CR: 1.e-2\n    eps: 0.7\n    U: 0.1\n    V: {V}\n    flip: {flip}\n"
_ADAPT = "numerics:\n    CFL: {CFL}\n    adaptive: 1\n    tol: 1e-8\n    dt: 1e-10\n    max_it: 10000\n    MC_order: {mc}\n"
_FIXED = "numerics:\n    tol: 1e-9\n    dt: 1e-10\n    max_it: 200\n"


def _grid(Nx, Ny, dx, dy, bc):
    return f"grid:\n    dx: {dx}\n    dy: {dy}\n    Nx: {Nx}\n    Ny: {Ny}\n" + bc


_BWR = """
properties:
    shear: 2.15
    bulk: 0.
    EOS: BWR
    T: 1.0
    rho0: 0.8
"""

# name -> dict(yaml=..., seed=..., extra=bool)
STEP_CASES = {
    # BASELINE config 1 as README.md prints it (D,N,N in x, fixed dt) ...
    'journal1d_readme': dict(yaml=_yaml(_grid(100, 1, '1.e-5', '1.', _XDN), _JOURNAL.format(V='0.', flip=False), _FIXED)),
    # ... and as the file that ships (periodic, adaptive)
    'journal1d_periodic': dict(yaml=_yaml(_grid(100, 1, '1.e-5', '1.', _PER), _JOURNAL.format(V='0.', flip=False),
                                          _ADAPT.format(CFL=0.25, mc=1))),
    # tests/test_flip_axes.py problem, small
    'journal2d_periodic': dict(yaml=_yaml(_grid(37, 29, '1.e-4', '1.e-4', _PER), _JOURNAL.format(V='0.', flip=False),
                                          _ADAPT.format(CFL=0.5, mc=1))),
    'journal2d_flip_mc0': dict(yaml=_yaml(_grid(31, 31, '1.e-4', '1.e-4', _PER), _JOURNAL.format(V='0.1', flip=True).replace('U: 0.1', 'U: 0.'),
                                          _ADAPT.format(CFL=0.5, mc=0))),
    'journal2d_mcm1_bulk': dict(yaml=_yaml(_grid(20, 70, '2.e-4', '1.e-4', _PER), _JOURNAL.format(V='-0.03', flip=False),
                                           _ADAPT.format(CFL=0.4, mc=-1), bulk=0.013)),
    # BASELINE config 2 (parabolic slider, D,N,N in x, periodic y), small
    'slider2d_xDN': dict(yaml=_yaml("grid:\n    Lx: 0.0762\n    Ly: 0.0762\n    Nx: 33\n    Ny: 65\n" + _XDN,
                                    "geometry:\n    type: parabolic\n    hmin: 2.54e-5\n    hmax: 5.08e-5\n    U: 4.57\n    V: 0.\n",
                                    _ADAPT.format(CFL=0.45, mc=1))),
    # Dirichlet/Neumann on both axes, different targets per side, slip-length field
    'slider2d_allDN_slip': dict(yaml=_yaml(_grid(18, 21, '2.e-5', '1.e-5',
                                                 "    xE: ['D', 'N', 'N']\n    xW: ['D', 'N', 'N']\n    yS: ['N', 'D', 'N']\n    yN: ['N', 'D', 'N']\n"
                                                 "    xE_D: 880.\n    xW_D: 875.\n    yS_D: 40.\n    yN_D: 45.\n"),
                                           "geometry:\n    type: inclined\n    hmin: 1.e-5\n    hmax: 2.e-5\n    U: 0.1\n    V: 0.02\n",
                                           _ADAPT.format(CFL=0.3, mc=0), bulk=0.02), extra=True),
    # LJ units, BWR EOS (mock-GP configs' fixed-form law), slip-length field
    'bwr2d_slip': dict(yaml=_yaml(_grid(16, 12, '1.', '1.', _PER),
                                  "geometry:\n    type: inclined\n    hmin: 11.\n    hmax: 12.\n    U: 0.12\n    V: 0.\n",
                                  "numerics:\n    CFL: 0.4\n    adaptive: 1\n    tol: 1e-8\n    dt: 0.05\n    max_it: 1000\n",
                                  prop=_BWR), extra=True, rel_rho=1e-3),
}

# Different masks on OPPOSITE sides: the reference assigns ghost row 0 through the E masks but
# values it through the W masks and the W target, ghost row -1 the other way round, likewise S / N
# (problem.py:682-707 vs :736-754), so the ghost value of one component is built from ANOTHER component of the
# adjacent cell.  Such conditions are not physical: U = V = 2 m/s puts rho, jx and jy on the same scale and a soft
# fluid (C1 = 3.5e4) with 1e-6 perturbations keeps the runs valid for the compared steps.  Legal
# combinations have the same number of D components on opposite sides (anything else raises in NumPy).
def _asym(xE, xW, yS, yN, vals):
    bc = f"    xE: {xE}\n    xW: {xW}\n    yS: {yS}\n    yN: {yN}\n"
    return bc + "".join(f"    {k}: {v}\n" for k, v in vals.items())


_P3, _SLIDER_U2 = "['P', 'P', 'P']", "geometry:\n    type: inclined\n    hmin: 1.e-5\n    hmax: 2.e-5\n    U: {U}\n    V: {V}\n"
STEP_CASES.update({
    # x: E = D,N,N / W = N,D,N  -> ghost row 0: rho <- 2 xW_D - jx, jx <- rho, jy <- jy; ghost row -1: rho <- jx, jx <- 2 xE_D - rho
    'asym_x_mc1': dict(yaml=_yaml(_grid(19, 23, '2.e-5', '1.e-5', _asym("['D', 'N', 'N']", "['N', 'D', 'N']", _P3, _P3,
                                                                       dict(xE_D=877.7007, xW_D=877.7107))),
                                  _SLIDER_U2.format(U=2.0, V=2.0), _ADAPT.format(CFL=0.3, mc=1), C1='3.5e4'), rel_j=1e-6),
    # y: S = N,N,D / N = D,N,N with MacCormack order -1 and a slip-length field
    'asym_y_mcm1_slip': dict(yaml=_yaml(_grid(17, 70, '2.e-5', '1.e-5', _asym(_P3, _P3, "['N', 'N', 'D']", "['D', 'N', 'N']",
                                                                             dict(yS_D=877.6907, yN_D=877.7057))),
                                        _SLIDER_U2.format(U=2.0, V=2.0), _ADAPT.format(CFL=0.3, mc=-1), bulk=0.01, C1='3.5e4'), extra=True, rel_j=1e-6),
    # both axes asymmetric (corner ghosts chain two cross-lookups), two D components per side, alternating order
    'asym_xy_mc0': dict(yaml=_yaml(_grid(21, 18, '2.e-5', '2.e-5', _asym("['D', 'D', 'N']", "['N', 'D', 'D']", "['D', 'N', 'D']", "['D', 'D', 'N']",
                                                                        dict(xE_D=877.7007, xW_D=877.7107, yS_D=877.6957, yN_D=877.7057))),
                                   _SLIDER_U2.format(U=2.0, V=2.0), _ADAPT.format(CFL=0.25, mc=0), C1='3.5e4'), rel_j=1e-6),
    # symmetric masks, non-periodic on both axes, with the two other MacCormack orders (so far only periodic cases used them)
    'slider2d_allDN_mcm1': dict(yaml=_yaml(_grid(18, 21, '2.e-5', '1.e-5',
                                                 "    xE: ['D', 'N', 'N']\n    xW: ['D', 'N', 'N']\n    yS: ['N', 'D', 'N']\n    yN: ['N', 'D', 'N']\n"
                                                 "    xE_D: 880.\n    xW_D: 875.\n    yS_D: 40.\n    yN_D: 45.\n"),
                                           "geometry:\n    type: inclined\n    hmin: 1.e-5\n    hmax: 2.e-5\n    U: 0.1\n    V: 0.02\n",
                                           _ADAPT.format(CFL=0.3, mc=-1), bulk=0.02)),
    'slider2d_xDN_mc0': dict(yaml=_yaml("grid:\n    Lx: 0.0762\n    Ly: 0.0762\n    Nx: 33\n    Ny: 65\n" + _XDN,
                                        "geometry:\n    type: parabolic\n    hmin: 2.54e-5\n    hmax: 5.08e-5\n    U: 4.57\n    V: 0.\n",
                                        _ADAPT.format(CFL=0.45, mc=0))),
})
ASYM_CASES = ['asym_x_mc1', 'asym_y_mcm1_slip', 'asym_xy_mc0']


_BAYADA = """
properties:
    EOS: Bayada
    rho0: 850.
    shear: 0.039
    bulk: 0.
    c_l: 1600.
    c_v: 352.
    piezo:
        name: {name}
        eta_v: 3.9e-5
        rho_l: 850.
        rho_v: 0.019
"""
_SLIDER = "geometry:\n    type: parabolic\n    hmin: 2.54e-5\n    hmax: 5.08e-5\n    U: 4.57\n    V: {V}\n"
_XDN850 = _XDN.replace('877.7007', '850.')

# viscosity modifiers (stress.py:305-326): piezo / thinning blocks make the viscosity a field
STEP_CASES.update({
    # examples/config/parabolic_1d_cav_bayada.yaml (density-based mixture viscosity, no pressure gradient needed)
    'cav1d_dukler': dict(yaml=_yaml("grid:\n    Lx: 0.0762\n    Ly: 1.\n    Nx: 100\n    Ny: 1\n" + _XDN850, _SLIDER.format(V='0.'),
                                    _ADAPT.format(CFL=0.45, mc=1), prop=_BAYADA.format(name='Dukler')), rel_rho=1e-3),
    'cav2d_mcadams': dict(yaml=_yaml("grid:\n    Lx: 0.0762\n    Ly: 0.0762\n    Nx: 21\n    Ny: 70\n" + _XDN850, _SLIDER.format(V='0.3'),
                                     _ADAPT.format(CFL=0.4, mc=0), prop=_BAYADA.format(name='McAdams')), rel_rho=1e-3),
    # pressure-based laws on the BASELINE config 2 slider: Barus alone; Barus + Eyring; Roelands + Carreau (+ slip
    # length, bulk viscosity, V != 0); Eyring alone
    'slider2d_barus': dict(yaml=_yaml("grid:\n    Lx: 0.0762\n    Ly: 0.0762\n    Nx: 33\n    Ny: 65\n" + _XDN, _SLIDER.format(V='0.'),
                                      _ADAPT.format(CFL=0.45, mc=1),
                                      prop=_DH.format(bulk=0., C1='3.5e10') + "    piezo:\n        name: Barus\n        aB: 2.e-8\n")),
    'slider2d_barus_eyring': dict(yaml=_yaml("grid:\n    Lx: 0.0762\n    Ly: 0.0762\n    Nx: 33\n    Ny: 65\n" + _XDN, _SLIDER.format(V='0.'),
                                             _ADAPT.format(CFL=0.45, mc=1),
                                             prop=_DH.format(bulk=0., C1='3.5e10') + "    piezo:\n        name: Barus\n        aB: 2.e-8\n"
                                             "    thinning:\n        name: Eyring\n        tauE: 5.e5\n")),
    'journal2d_roelands_carreau_slip': dict(yaml=_yaml(_grid(20, 70, '2.e-4', '1.e-4', _PER), _JOURNAL.format(V='-0.03', flip=False),
                                                       _ADAPT.format(CFL=0.4, mc=-1),
                                                       prop=_DH.format(bulk=0.013, C1='3.5e10') +
                                                       "    piezo:\n        name: Roelands\n        mu_inf: 1.e-3\n        p_ref: 1.96e8\n        z: 0.68\n"
                                                       "    thinning:\n        name: Carreau\n        mu_inf: 1.e-3\n        lam: 0.02\n        a: 2.\n        N: 0.8\n"),
                                            extra=True),
    'journal1d_eyring': dict(yaml=_yaml(_grid(100, 1, '1.e-5', '1.', _XDN), _JOURNAL.format(V='0.', flip=False), _FIXED,
                                        prop=_DH.format(bulk=0., C1='3.5e10') + "    thinning:\n        name: Eyring\n        tauE: 5.e5\n")),
})
VISC_CASES = ['cav1d_dukler', 'cav2d_mcadams', 'slider2d_barus', 'slider2d_barus_eyring',
              'journal2d_roelands_carreau_slip', 'journal1d_eyring']


def perturb(q, rho0, U, seed, rel_rho=1e-6, rel_j=1e-3):
    """Seeded perturbation of SURVEY 8(d): q0 *= 1 + 1e-6 N(0,1) (1e-4 makes |v| ~ 1e5 m/s with C1=3.5e10); q1, q2 += 1e-3 rho0 U N(0,1)."""
    rng = np.random.default_rng(seed)
    shape = q.shape[1:]
    q[0] *= 1 + rel_rho * rng.standard_normal(shape)
    q[1] += rel_j * rho0 * U * rng.standard_normal(shape)
    q[2] += rel_j * rho0 * U * rng.standard_normal(shape)


def slip_field(shape, scale, seed=7):
    rng = np.random.default_rng(seed)
    return scale * rng.random(shape)


def inputs(name):
    """Sanitised dicts + extra field of a named case."""
    case = STEP_CASES[name]
    d = config.read_yaml_string(case['yaml'])
    grid = d['grid']
    shape = (grid['Nx'] + 2, grid['Ny'] + 2)
    extra = None
    if case.get('extra'):
        hscale = 1e-6 if d['properties']['EOS'] == 'DH' else 1.0
        extra = slip_field(shape, hscale)
    return d, extra


def build_case(name, cls=None, seed=1234):
    """CPU oracle for the case, perturbed, ghosts filled, dt re-initialised as _pre_run does."""
    from oracle.fv import FVOracle
    cls = cls or FVOracle
    d, extra = inputs(name)
    o = cls(d['grid'], d['geometry'], d['properties'], d['numerics'], extra=extra)
    Umag = max(abs(d['geometry']['U']), abs(d['geometry']['V']))
    perturb(o.q, d['properties']['rho0'], Umag, seed, STEP_CASES[name].get('rel_rho', 1e-6), STEP_CASES[name].get('rel_j', 1e-3))
    o.fill_ghosts()
    if d['numerics']['adaptive']:
        o.dt = d['numerics']['CFL'] * o.dt_crit
    return o
