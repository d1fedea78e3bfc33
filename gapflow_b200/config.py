"""YAML input schema of GaPFlow, unchanged (reference: GaPFlow/io.py:100-452).

Same sections (options, grid, geometry, numerics, properties, gp, db, md), same keys,
same defaults and the same quirks (``init_seed`` reads ``init_width``, io.py:439; the
journal geometry test ``'CR' and 'eps' in keys``, io.py:240), so a reference YAML file
yields the same sanitised dictionaries.  Table-driven instead of the reference's
hand-written blocks.
"""
import io as _io

import yaml

_EOS_TABLE = {
    'DH': (['rho0', 'P0', 'C1', 'C2'], [877.7007, 101325, 3.5e10, 1.23]),
    'PL': (['rho0', 'P0', 'alpha'], [1.1853, 101325, 0.]),
    'vdW': (['M', 'T', 'a', 'b'], [39.948, 100., 1.355, 0.03201]),
    'MT': (['rho0', 'P0', 'K', 'n'], [700., 0.101e6, .557e9, 7.33]),
    'cubic': (['a', 'b', 'c', 'd'], [15.2, -9.6, 3.35, -0.07]),
    'BWR': (['T', 'gamma'], [2., 3.0]),
    'Bayada': (['rho_l', 'rho_v', 'c_l', 'c_v'], [850., 0.019, 1600., 352.]),
    'MD': (['rho0'], [1.]),
}
_PIEZO = {'Roelands': (['mu_inf', 'p_ref', 'z'], [1.e-3, 1.96e8, 0.68]),
          'Barus': (['aB'], [20e-9]),
          'Dukler': (['eta_v', 'rho_l', 'rho_v'], [3.9e-5, 850., 0.019]),
          'McAdams': (['eta_v', 'rho_l', 'rho_v'], [3.9e-5, 850., 0.019])}
_THIN = {'Carreau': (['mu_inf', 'lam', 'a', 'N'], [1.e-9, 1e-6, 2., 0.6]),
         'Eyring': (['tauE'], [5.e5])}
_GEO_KEYS = {'inclined': ['hmax', 'hmin'], 'parabolic': ['hmin', 'hmax'], 'cdc': ['hmin', 'hmax', 'b'],
             'asperity': ['hmin', 'hmax']}


def sanitize_options(d):
    return {'output': str(d.get('output', 'example')), 'write_freq': int(d.get('write_freq', 1000)),
            'use_tstamp': bool(d.get('use_tstamp', True)), 'silent': bool(d.get('silent', False))}


def _axis(out, d, n, L, dl, n_default):
    out[n] = int(d.get(n, n_default))
    if L in d:
        out[L] = float(d.get(L, 1.))
        out[dl] = out[L] / out[n]
    elif dl in d:
        out[dl] = float(d.get(dl, 0.1))
        out[L] = out[dl] * out[n]
    else:
        raise IOError(f"Must specify grid size ({n}) with either {dl} or {L}.")


def sanitize_grid(d):
    out = {}
    _axis(out, d, 'Nx', 'Lx', 'dx', 100)
    _axis(out, d, 'Ny', 'Ly', 'dy', 1)
    out['dim'] = int(out['Nx'] > 1) + int(out['Ny'] > 1)
    for a, b, x_axis in (('xE', 'xW', True), ('yS', 'yN', False)):
        for side in (a, b):
            bc = list(d.get(side, ['P', 'P', 'P']))
            assert all(t in ['P', 'N', 'D'] for t in bc)
            for kind in 'PDN':
                out[f'bc_{side}_{kind}'] = [t == kind for t in bc]
        for side in (a, b):
            if any(out[f'bc_{side}_D']):
                # io.py:189-197: x sides default to 1., y sides to None (then an error)
                out[f'bc_{side}_D_val'] = d.get(f'{side}_D', 1. if x_axis else None)
                if out[f'bc_{side}_D_val'] is None:
                    raise IOError("Need to specify Dirichlet BC value")
        assert all(e == w for e, w in zip(out[f'bc_{a}_P'], out[f'bc_{b}_P']))
    return out


def sanitize_geometry(d):
    out = {'U': float(d.get('U', 1.)), 'V': float(d.get('V', 0.)), 'type': str(d.get('type', 'none')),
           'flip': bool(d.get('flip', False))}
    t = out['type']
    if t not in ['journal', 'inclined', 'parabolic', 'cdc', 'asperity']:
        raise IOError("Specify a valid geometry type")
    if t == 'journal':
        if 'eps' in d.keys():            # reference: `"CR" and 'eps' in d.keys()`
            out['CR'] = float(d.get('CR'))
            out['eps'] = float(d.get('eps'))
        elif 'hmax' in d.keys():         # reference: `"hmin" and 'hmax' in d.keys()`
            out['hmin'] = float(d.get('hmin'))
            out['hmax'] = float(d.get('hmax'))
        else:
            raise IOError("Need to specify either clearance ratio and eccentrity or min/max gap height")
    else:
        for k in _GEO_KEYS[t]:
            out[k] = float(d.get(k))
        if t == 'asperity':
            out['num'] = int(d.get('num', 1))
    return out


def sanitize_properties(d):
    out = {'shear': float(d.get('shear', -1.))}
    if out['shear'] < 0.:
        raise IOError("Specify a a (non-negative) shear viscosity")
    out['bulk'] = float(d.get('bulk', -1.))
    out['EOS'] = str(d.get('EOS', 'none'))
    if out['EOS'] not in _EOS_TABLE:
        raise IOError("Specify a valid equation of state")
    keys, defaults = _EOS_TABLE[out['EOS']]
    for k, de in zip(keys, defaults):
        out[k] = float(d.get(k, de))
    if 'rho0' not in out:
        out['rho0'] = float(d.get('rho0', 1.))
    for sec, table in (('piezo', _PIEZO), ('thinning', _THIN)):
        if sec in d.keys():
            out[sec] = {'name': str(d[sec].get('name', 'none'))}
            if out[sec]['name'] in table:
                keys, defaults = table[out[sec]['name']]
                for k, de in zip(keys, defaults):
                    out[sec][k] = float(d[sec].get(k, de))
    if 'elastic' in d.keys():
        e = d['elastic']
        out['elastic'] = {'enabled': True, 'E': float(e.get('E', 210e09)), 'v': float(e.get('v', 0.3)),
                          'alpha_underrelax': float(e.get('alpha_underrelax', 1e-03)),
                          'n_images': int(e.get('n_images', 10))}
    else:
        out['elastic'] = {'enabled': False}
    return out


def sanitize_numerics(d):
    out = {'tol': float(d.get('tol', 1e-6)), 'max_it': int(d.get('max_it', 1000)),
           'dt': float(d.get('dt', 3e-10)), 'adaptive': bool(d.get('adaptive', False)),
           'CFL': float(d.get('CFL', 0.5)), 'MC_order': int(d.get('MC_order', 1))}
    # the one key the reference does not have: the arithmetic contract of the pass kernels (include/gapflow_b200.h:
    # gpf_fv_cfg.arith).  Absent = 'exact' = the reference's NumPy roundings; 'fast' = FMA contraction, 1e-12 per step
    if 'arith' in d:
        out['arith'] = str(d['arith'])
        if out['arith'] not in ('exact', 'fast'):
            raise IOError("numerics.arith must be 'exact' or 'fast'")
    return out


def sanitize_gp(d):
    out = {'press_gp': 'press' in d.keys(), 'shear_gp': 'shear' in d.keys()}
    for sk in ('press', 'shear'):
        if sk not in d.keys():
            continue
        ds = d[sk]
        o = {'atol': float(ds.get('atol', 1.)), 'rtol': float(ds.get('rtol', 0.5)),
             'obs_stddev': float(ds.get('obs_stddev', 0.)), 'fix_noise': bool(ds.get('fix_noise', True)),
             'max_steps': int(ds.get('max_steps', 5)), 'pause_steps': int(ds.get('pause_steps', 100)),
             'active_learning': bool(ds.get('active_learning', True))}
        if sk == 'press':
            o['active_dims'] = list(ds.get('active_dims', [0, 3]))
        else:
            ad = ds.get('active_dims', {})
            o['active_dims_x'] = ad.get('x', [0, 1, 3])
            o['active_dims_y'] = ad.get('y', [0, 2, 3])
        out[sk] = o
    return out


def sanitize_db(d):
    out = {'dtool_path': d.get('dtool_path', None), 'init_size': int(d.get('init_size', 5)),
           'init_method': str(d.get('init_method', 'lhc')), 'init_width': float(d.get('init_width', 1e-2)),
           'init_seed': int(d.get('init_width', 123))}     # sic: io.py:439 reads 'init_width'
    assert out['init_method'] in ['rand', 'lhc', 'sobol']
    return out


def sanitize_md(d):
    return d


_SANITIZERS = {'options': sanitize_options, 'grid': sanitize_grid, 'geometry': sanitize_geometry,
               'numerics': sanitize_numerics, 'properties': sanitize_properties, 'gp': sanitize_gp,
               'db': sanitize_db, 'md': sanitize_md}


def read_yaml_input(file):
    """io.py:100-125: parse a YAML stream into the eight sanitised sections (None if absent)."""
    raw = yaml.full_load(file)
    return {k: (f(raw[k]) if raw.get(k) is not None else None) for k, f in _SANITIZERS.items()}


def read_yaml_string(s):
    with _io.StringIO(s) as f:
        return read_yaml_input(f)
