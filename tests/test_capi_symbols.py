"""CPU: libgapflow_b200.so builds for sm_100a, loads, and exports every symbol that
include/gapflow_b200.h declares (no compute calls: there is no GPU here)."""
import ctypes
import os
import re

import pytest

from gapflow_b200 import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "gapflow_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(gpf_[a-z0-9_]+)\s*\(", txt)))


@pytest.fixture(scope="module")
def lib():
    from gapflow_b200 import build
    build.build()
    return _capi.load()


def test_every_declared_symbol_is_exported_and_bound(lib):
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
        assert n in _capi.SIGNATURES, f"{n} declared in the header but not bound in _capi.SIGNATURES"
    for n in _capi.SIGNATURES:
        assert n in names, f"{n} bound but not declared in the header"


def test_struct_layout_matches_header(lib):
    # sizes implied by the header's field lists (natural alignment)
    assert ctypes.sizeof(_capi.FvCfg) == 8 + 16 + 48 + 48 + 32 + 32 + 8 + 64 + 8 + 8 + 24 + 8 + 4 + 28
    assert ctypes.sizeof(_capi.FvScalars) == 8 + 24 + 40 + 8 + 48 + 16


def test_bc_table_reproduces_reference_cross_lookup(lib):
    from gapflow_b200 import config
    from tests.cases import STEP_CASES
    d = config.read_yaml_string(STEP_CASES['slider2d_allDN_slip']['yaml'])
    c = _capi.make_cfg(d['grid'], d['geometry'], d['properties'], d['numerics'])
    # ghost x0 takes the W target, x1 the E target; y0 the N target, y1 the S target (problem.py:736-754)
    assert list(c.bc_val) == [875., 880., 45., 40.]
    assert [list(r) for r in c.bc_type] == [[1, 2, 2], [1, 2, 2], [2, 1, 2], [2, 1, 2]]
    assert [list(r) for r in c.bc_src] == [[0, 1, 2]] * 4


def test_invalid_configs_are_rejected(lib):
    from gapflow_b200 import config
    from tests.cases import STEP_CASES
    d = config.read_yaml_string(STEP_CASES['journal2d_periodic']['yaml'])
    g = dict(d['grid'])
    g['bc_xE_P'] = [True, False, False]
    g['bc_xW_P'] = [True, False, False]
    g['bc_xE_D'] = [False, True, True]
    g['bc_xW_D'] = [False, True, True]
    g['bc_xE_D_val'] = g['bc_xW_D_val'] = 1.0
    with pytest.raises(_capi.GapflowError):
        _capi.make_cfg(g, d['geometry'], d['properties'], d['numerics'])
    p = dict(d['properties'])
    p['EOS'] = 'MD'
    with pytest.raises(_capi.GapflowError):
        _capi.make_cfg(d['grid'], d['geometry'], p, d['numerics'])


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_capi, "_lib", None)
    monkeypatch.setattr(_capi, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_capi.GapflowLibraryError):
        _capi.load()
