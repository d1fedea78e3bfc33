"""CPU, where the reference tree is present: the YAML front end (gapflow_b200/config.py) against the reference's own
sanitisers (GaPFlow/io.py:100-460) -- same dicts, same types, for the reference's example configurations and for every
YAML text the tests and the benchmark of this repository use.  io.py imports polars for its CSV writer only; the module
is loaded by file path with an empty stand-in for that import."""
import contextlib
import glob
import io
import os
import types

import numpy as np
import pytest

from gapflow_b200 import config
from oracle import ref_loader

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")


@pytest.fixture(scope="module")
def ref_io():
    return ref_loader.load_with_stand_ins("GaPFlow/io.py", "_ref_io", {'polars': types.ModuleType('polars')})


def _norm(x):
    if isinstance(x, dict):
        return {k: _norm(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return [_norm(v) for v in x]
    if isinstance(x, np.ndarray):
        return _norm(x.tolist())
    if isinstance(x, np.floating):
        return float(x)
    if isinstance(x, np.integer):
        return int(x)
    if isinstance(x, np.bool_):
        return bool(x)
    return x


def _diff(a, b, path=""):
    if isinstance(a, dict) and isinstance(b, dict):
        out = []
        for k in sorted(set(a) | set(b), key=str):
            if k not in a or k not in b:
                out.append(f"{path}/{k}: only in {'ours' if k not in a else 'the reference'}")
            else:
                out += _diff(a[k], b[k], f"{path}/{k}")
        return out
    if type(a) is not type(b) or a != b:
        return [f"{path}: reference {a!r} ({type(a).__name__}), ours {b!r} ({type(b).__name__})"]
    return []


def _both(ref_io, text):
    res = []
    for mod in (ref_io, config):
        with io.StringIO(text) as fh, contextlib.redirect_stdout(io.StringIO()):
            try:
                res.append(_norm(mod.read_yaml_input(fh)))
            except Exception as e:          # noqa: BLE001  (both must refuse the same inputs)
                res.append(type(e).__name__)
    return res


def _texts():
    from tests import cases, surrogate_cases as SC
    from tests.test_problem_gpu import DECAY, JOURNAL2D, SOMMERFELD
    import bench
    out = {f"example:{os.path.basename(f)}": open(f).read()
           for f in sorted(glob.glob(os.path.join(ref_loader.REF_ROOT, "examples", "config", "*.yaml")))}
    out.update({f"case:{k}": v['yaml'] for k, v in cases.STEP_CASES.items()})
    out.update({"sommerfeld": SOMMERFELD, "journal2d": JOURNAL2D.format(n=50), "decay": DECAY})
    out["slip1d"] = SC.inputs('slip1d')[1]
    out["slip2d"] = SC.inputs('slip2d', nx=24, ny=20)[1]
    out["bench:journal"] = bench.JOURNAL_YAML.format(Nx=8192, Ny=8192, U="0.1", V="0.", flip=False)
    out["bench:slider"] = bench.SLIDER_YAML.format(Nx=4096, Ny=4096, Lx=0.0762)
    return out


def test_sanitised_inputs_equal_the_reference_s(ref_io):
    texts = _texts()
    assert len(texts) > 30
    bad = {}
    for name, text in texts.items():
        ref, ours = _both(ref_io, text)
        if isinstance(ref, str) or isinstance(ours, str):
            if ref != ours:
                bad[name] = [f"reference: {ref if isinstance(ref, str) else 'accepted'}, ours: {ours if isinstance(ours, str) else 'accepted'}"]
            continue
        d = _diff(ref, ours)
        if d:
            bad[name] = d
    assert not bad, bad


def _mutations():
    pop = lambda sec, key: (lambda d: d[sec].pop(key, None))            # noqa: E731
    upd = lambda sec, **kw: (lambda d: d[sec].update(**kw))              # noqa: E731
    fv = {
        'no_grid': lambda d: d.pop('grid'), 'no_numerics': lambda d: d.pop('numerics'), 'no_options': lambda d: d.pop('options'),
        'no_Nx': pop('grid', 'Nx'), 'no_dx_no_Lx': lambda d: [d['grid'].pop(k, None) for k in ('dx', 'Lx')],
        'both_dx_Lx': upd('grid', dx=1e-5, Lx=1.0), 'D_without_value': pop('grid', 'xE_D'),
        'bad_bc_letter': upd('grid', xE=['X', 'N', 'N']), 'P_on_one_side': upd('grid', xE=['P', 'P', 'P']),
        'str_Nx': upd('grid', Nx='100'), 'float_Nx': upd('grid', Nx=100.0),
        'unknown_eos': upd('properties', EOS='foo'), 'no_eos': pop('properties', 'EOS'), 'no_shear': pop('properties', 'shear'),
        'unknown_geo': upd('geometry', type='foo'), 'no_geo_type': pop('geometry', 'type'),
        'mc_order_2': upd('numerics', MC_order=2),
        'piezo_unknown': upd('properties', piezo={'name': 'foo'}),
        'thinning_eyring': upd('properties', thinning={'name': 'Eyring', 'tauE': 5e5}),
        'thinning_missing_parameter': upd('properties', thinning={'name': 'Eyring'}),
    }
    gp = {
        'gp_no_press': pop('gp', 'press'), 'gp_no_shear': pop('gp', 'shear'), 'gp_empty': lambda d: d.update(gp={}),
        'no_db': lambda d: d.pop('db', None), 'db_unknown_method': upd('db', init_method='foo'),
        'gp_press_atol_missing': lambda d: d['gp']['press'].pop('atol', None),
    }
    return fv, gp


def test_malformed_inputs_are_refused_or_defaulted_like_the_reference(ref_io):
    """io.py's error behaviour: the same inputs raise the same exception type, the same inputs are accepted with the same
    defaults filled in."""
    import copy
    import yaml
    texts = _texts()
    fv, gp = _mutations()
    bad = {}
    for base, muts in ((yaml.safe_load(texts['case:slider2d_xDN']), fv), (yaml.safe_load(texts['slip1d']), gp)):
        for name, f in muts.items():
            d = copy.deepcopy(base)
            f(d)
            ref, ours = _both(ref_io, yaml.safe_dump(d))
            if isinstance(ref, str) or isinstance(ours, str):
                if ref != ours:
                    bad[name] = (ref if isinstance(ref, str) else 'accepted', ours if isinstance(ours, str) else 'accepted')
            elif _diff(ref, ours):
                bad[name] = _diff(ref, ours)
    assert not bad, bad
