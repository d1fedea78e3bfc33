"""CPU, where the reference tree is present: the gap-height fields (gapflow_b200/topography.py, oracle/fv.py) against the
reference's own Topography class (GaPFlow/topography.py:38-255), bit for bit, for every geometry type, with and without
`flip`, 1D and 2D grids, and for the row windows the slab decomposition asks for.  topography.py imports
ContactMechanics (elastic deformation, out of scope) and muGrid's field wrapper; the module is loaded by file path with
stand-ins for both: the wrapper becomes a plain array holder, so Topography.__init__ itself runs unchanged."""
import types

import numpy as np
import pytest

from gapflow_b200 import config
from gapflow_b200.topography import make_topography, midpoint_grid
from oracle import fv as OF
from oracle import ref_loader

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")


class _Holder:
    def __init__(self, shape):
        self.p = np.zeros(shape)


@pytest.fixture(scope="module")
def ref_topo():
    mg, mf = types.ModuleType('muGrid'), types.ModuleType('muGrid.Field')
    mf.wrap_field = lambda f: f
    mg.Field = mf
    return ref_loader.load_with_stand_ins("GaPFlow/topography.py", "_ref_topography",
                                          {'ContactMechanics': types.ModuleType('ContactMechanics'), 'muGrid': mg, 'muGrid.Field': mf})


class _Fields:
    def __init__(self, shape):
        self.f = {'topography': _Holder((4, *shape)), 'x': _Holder(shape), 'y': _Holder(shape), 'pressure': _Holder(shape)}

    def get_real_field(self, name):
        return self.f[name]


GRID = "grid:\n    Lx: {Lx}\n    Ly: {Ly}\n    Nx: {Nx}\n    Ny: {Ny}\n    xE: ['P', 'P', 'P']\n    xW: ['P', 'P', 'P']\n    yS: ['P', 'P', 'P']\n    yN: ['P', 'P', 'P']\n"
TAIL = "numerics:\n    dt: 1e-10\nproperties:\n    shear: 1.\n    bulk: 0.\n    EOS: DH\n"
GEOS = {
    'journal_cr': "    type: journal\n    CR: 1.e-2\n    eps: 0.7\n",
    'journal_h': "    type: journal\n    hmin: 1.e-6\n    hmax: 3.e-6\n",
    'inclined': "    type: inclined\n    hmin: 2.e-6\n    hmax: 5.e-6\n",
    'parabolic': "    type: parabolic\n    hmin: 2.54e-5\n    hmax: 5.08e-5\n",
    'cdc': "    type: cdc\n    hmin: 1.e-6\n    hmax: 2.e-6\n    b: 1.e-4\n",
    'asperity': "    type: asperity\n    hmin: 1.e-6\n    hmax: 2.e-6\n    num: 1\n",
}


def _inputs(geo, Nx, Ny, flip):
    text = GRID.format(Lx=1.3e-3, Ly=0.7e-3 if Ny > 1 else 1., Nx=Nx, Ny=Ny) + "geometry:\n" + GEOS[geo] + \
        f"    U: 0.1\n    V: 0.\n    flip: {flip}\n" + TAIL
    return config.read_yaml_string(text)


@pytest.mark.parametrize('geo', list(GEOS))
@pytest.mark.parametrize('Nx, Ny, flip', [(37, 1, False), (29, 23, False), (24, 24, True)])
def test_gap_height_equals_the_reference_s(ref_topo, geo, Nx, Ny, flip):
    d = _inputs(geo, Nx, Ny, flip)
    g = d['grid']
    shape = (g['Nx'] + 2, g['Ny'] + 2)
    T = ref_topo.Topography(_Fields(shape), g, d['geometry'], d['properties'])
    ours = make_topography(g, d['geometry'])
    assert np.array_equal(ours[:3], T.full[:3]), geo
    assert np.array_equal(OF.make_topography(g, d['geometry'])[:3], T.full[:3]), geo
    xx, yy = midpoint_grid(g)
    assert np.array_equal(xx, T.x) and np.array_equal(yy, T.y)
    if geo != 'asperity' and not flip:
        # a slab's rows are the rows of the global field (dist.py, bench.py: every rank builds only its own)
        for a, b in ((0, 9), (8, shape[0]), (5, 6)):
            assert np.array_equal(make_topography(g, d['geometry'], rows=(a, b))[:3], T.full[:3, a:b]), (geo, a, b)
