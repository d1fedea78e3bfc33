"""CPU: host-side logic of the binding that needs no device."""
import pytest


def test_viscosity_modifiers_are_refused_not_ignored():
    """A 'piezo' / 'thinning' block (stress.py:306-326) changes the reference's result; the CUDA path does
    not implement it yet and must say so instead of running Newtonian."""
    import copy
    from gapflow_b200 import _capi
    from tests.cases import inputs
    d, _ = inputs('journal1d_readme')
    for sec, blk in (('piezo', {'name': 'Barus', 'aB': 2e-8}), ('thinning', {'name': 'Eyring', 'tauE': 5e5})):
        prop = copy.deepcopy(d['properties'])
        prop[sec] = blk
        try:
            _capi.make_cfg(d['grid'], d['geometry'], prop, d['numerics'])
        except NotImplementedError as e:
            assert sec in str(e)
        except _capi.GapflowLibraryError:
            pytest.skip("library not built")
        else:
            raise AssertionError(f"{sec} block silently accepted")
