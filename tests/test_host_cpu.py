"""CPU: host-side logic of the binding that needs no device."""
import copy

import pytest


def _cfg(prop, **kw):
    from gapflow_b200 import _capi
    from tests.cases import inputs
    d, _ = inputs('journal1d_readme')
    p = copy.deepcopy(d['properties'])
    p.update(prop)
    try:
        return _capi.make_cfg(d['grid'], d['geometry'], p, d['numerics'], **kw), p
    except _capi.GapflowLibraryError:
        pytest.skip("library not built")


def test_viscosity_blocks_select_the_field_path():
    """A 'piezo' / 'thinning' block (stress.py:305-326) makes the viscosity a field: cfg.visc_field is set
    (never silently Newtonian) and the law ids / parameter order are those of include/gapflow_b200.h."""
    from gapflow_b200 import _capi
    c, _ = _cfg({})
    assert c.visc_field == 0
    c, p = _cfg({'piezo': {'name': 'Barus', 'aB': 2e-8}})
    assert c.visc_field == 1
    pz, pp, th, tp = _capi.viscosity_args(p)
    assert (pz, list(pp), th, tp) == (2, [2e-8], 0, None)
    c, p = _cfg({'thinning': {'name': 'Carreau', 'mu_inf': 1e-3, 'lam': 0.02, 'a': 2., 'N': 0.8},
                 'piezo': {'name': 'Roelands', 'mu_inf': 1e-3, 'p_ref': 1.96e8, 'z': 0.68}})
    assert c.visc_field == 1
    pz, pp, th, tp = _capi.viscosity_args(p)
    assert (pz, list(pp), th, list(tp)) == (3, [1e-3, 1.96e8, 0.68], 3, [1e-3, 0.02, 2., 0.8])
    # a name the reference does not know: identity law, still the field path (viscosity.py:64, 94)
    c, p = _cfg({'piezo': {'name': 'none'}, 'thinning': {'name': 'Eyring', 'tauE': 5e5}})
    assert _capi.viscosity_args(p)[0] == 1 and _capi.viscosity_args(p)[2] == 2 and c.visc_field == 1
    for name, want in (('Dukler', 4), ('McAdams', 5)):
        _, p = _cfg({'piezo': {'name': name, 'eta_v': 3.9e-5, 'rho_l': 850., 'rho_v': 0.019}})
        pz, pp, _, _ = _capi.viscosity_args(p)
        assert pz == want and list(pp) == [3.9e-5, 850., 0.019]


def test_viscosity_blocks_are_refused_on_slabs():
    """The pressure gradient at a halo row needs a second halo row: refused, not silently wrong."""
    with pytest.raises(NotImplementedError):
        _cfg({'thinning': {'name': 'Eyring', 'tauE': 5e5}}, halo=(True, False))
