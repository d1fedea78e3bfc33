"""Runs the reference's own ``Problem`` class (GaPFlow/problem.py) for FIXED-FORM closures in this container --
TEST INFRASTRUCTURE ONLY, and only where /root/reference exists.

``import GaPFlow`` needs jax, tinygp, jaxopt, muGrid, ContactMechanics, polars, dtoolcore, netCDF4, matplotlib ...,
none of which is installed.  The fixed-form update loop (problem.py:509-768, models/stress.py, integrate.py,
models/pressure.py / viscous.py / sound.py / viscosity.py, topography.py) is plain NumPy; what it takes from those
packages is

* muGrid: a container of named arrays with a ``.p`` view (``GlobalFieldCollection.real_field``) -- replaced here by a
  dictionary of NumPy arrays of the same shapes (components first, then the padded grid);
* jax: decorators and type names at import time (``jax.jit``, ``jax.Array``); nothing of it executes without a GP.

So the missing packages are replaced by inert stand-ins (every function of theirs that would compute something
raises), GaPFlow's own md / db / viz sub-packages (MD runners, training database, plotting: outside this path) by empty
ones, and the reference's modules are imported unchanged from where they lie.  What runs is the reference's code:
``Problem.__init__``, ``_pre_run``, ``update``, the ghost-cell routines, the stress models' ``update``, ``Topography``.
tests/test_oracle_vs_reference_problem.py holds oracle/fv.py to it bit for bit.
"""
import importlib
import os
import sys
import types

import numpy as np

from .ref_loader import REF_ROOT, available  # noqa: F401

_PKG = "GaPFlow"


class _Inert:
    """Attribute / call sink for names that are imported but must never compute anything on this path."""

    def __init__(self, name):
        self._name = name

    def __getattr__(self, k):
        if k.startswith('__'):
            raise AttributeError(k)
        return _Inert(f"{self._name}.{k}")

    def __call__(self, *a, **k):
        raise RuntimeError(f"{self._name} is a stand-in (package not installed): the fixed-form path must not call it")

    def __getitem__(self, k):
        return self

    def __mro_entries__(self, bases):
        return (object,)

    def __or__(self, other):             # `Name | None` in annotations
        return self

    __ror__ = __or__


_installed = []


def _inert_module(name, **attrs):
    _installed.append(name)
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    def missing(k, _n=name):                                         # PEP 562
        if k.startswith('__'):
            raise AttributeError(k)
        return _Inert(f"{_n}.{k}")
    m.__getattr__ = missing
    m.__path__ = []                                                  # lets `import name.sub` find our sub-stubs
    sys.modules[name] = m
    return m


class _Field:
    def __init__(self, shape):
        self.p = np.zeros(shape)

    @property
    def s(self):
        return self.p


class _FieldCollection:
    """muGrid.GlobalFieldCollection as problem.py uses it: named real fields, components first."""

    def __init__(self, nb_grid_pts, *a, **k):
        self.nb_grid_pts = tuple(int(n) for n in nb_grid_pts)
        self.fields = {}

    def real_field(self, name, components=()):
        if name not in self.fields:
            self.fields[name] = _Field(tuple(components) + self.nb_grid_pts)
        return self.fields[name]

    register_real_field = real_field

    def get_real_field(self, name):
        return self.fields[name]


class FileIONetCDF:
    """muGrid.FileIONetCDF as problem.py:166-205, 629 uses it; frames are kept in memory (``opened`` lists the files)."""
    opened = []

    def __init__(self, path, mode=None, *a, **k):
        self.path, self.frames, self.closed = path, [], False
        FileIONetCDF.opened.append(self)

    def register_field_collection(self, fc, field_names=None):
        self.fc, self.names = fc, list(field_names if field_names is not None else fc.fields)

    def append_frame(self):
        return self

    def write(self):
        self.frames.append({n: self.fc.get_real_field(n).p.copy() for n in self.names})

    def close(self):
        self.closed = True


class DataFrame:
    """polars.DataFrame(dict).write_csv(path) as io.py:95-97 uses it: header = keys, one row per entry"""

    def __init__(self, data):
        self.data = {k: list(v) for k, v in data.items()}

    def write_csv(self, fname):
        import csv
        with open(fname, 'w', newline='') as f:
            w = csv.writer(f)
            w.writerow(list(self.data))
            w.writerows(zip(*[[float(x) if np.ndim(x) == 0 else np.asarray(x).tolist() for x in col] for col in self.data.values()]))


def _install_stand_ins():
    from . import tinygp_stand_in as T
    jit = lambda f=None, **k: f if f is not None else (lambda g: g)          # noqa: E731
    jax = _inert_module('jax', jit=jit, Array=np.ndarray, numpy=np, grad=T.grad, vmap=T.vmap,
                        config=types.SimpleNamespace(update=lambda *a, **k: None))
    sys.modules['jax.numpy'] = np
    _installed.append('jax.numpy')
    jax.random = _inert_module('jax.random', key=T.Random.key, split=T.Random.split, normal=T.Random.normal,
                               uniform=T.Random.uniform, choice=T.Random.choice)
    _inert_module('jaxopt', ScipyMinimize=T.ScipyMinimize)
    _inert_module('tinygp', GaussianProcess=T.GaussianProcess,
                  kernels=types.SimpleNamespace(stationary=types.SimpleNamespace(Matern32=T.Matern32),
                                                distance=types.SimpleNamespace(L2Distance=T.L2Distance)),
                  transforms=types.SimpleNamespace(Linear=T.Linear))
    mg = _inert_module('muGrid', GlobalFieldCollection=_FieldCollection, FileIONetCDF=FileIONetCDF,
                       OpenMode=types.SimpleNamespace(Overwrite='w', Read='r', Append='a'))
    mg.Field = _inert_module('muGrid.Field', wrap_field=lambda f: f)

    class YAML:                                  # ruamel.yaml.YAML(): configured at import time, used only for dtool readmes
        explicit_start = False

        def indent(self, *a, **k):
            pass
    for name in ('ContactMechanics', 'polars', 'dtoolcore', 'dtool_lookup_api', 'netCDF4', 'matplotlib', 'matplotlib.pyplot',
                 'matplotlib.animation', 'IPython', 'IPython.display', 'ruamel', 'ruamel.yaml'):
        if name not in sys.modules:
            _inert_module(name, **({'YAML': YAML} if name == 'ruamel.yaml' else {'DataFrame': DataFrame} if name == 'polars' else {}))


_ref = None


def reference_package():
    """namespace(Problem, Database, Mock, read_yaml_input): the reference's classes, imported from REF_ROOT with the
    stand-ins above.  ``Mock`` is the reference's mock MD with the dtool bookkeeping of ``MolecularDynamics.run``
    (md/base.py:190-240: dataset directories, readme files) left out -- ``run`` is build_input_files + read_output."""
    global _ref
    if _ref is not None:
        return _ref
    if not available():
        raise RuntimeError("reference tree not present")
    if _PKG in sys.modules:
        raise RuntimeError("a GaPFlow package is already imported")
    _install_stand_ins()
    root = os.path.join(REF_ROOT, _PKG)
    pkg = types.ModuleType(_PKG)
    pkg.__path__ = [root]
    pkg.__version__ = "reference"
    sys.modules[_PKG] = pkg
    _installed.append(_PKG)
    # the reference's own sub-packages that lie outside the path: plotting, MD runners (LAMMPS, MPI)
    viz = _inert_module(f'{_PKG}.viz')
    viz.plotting = _inert_module(f'{_PKG}.viz.plotting')
    viz.animations = _inert_module(f'{_PKG}.viz.animations')
    md = types.ModuleType(f'{_PKG}.md')
    md.__path__ = [os.path.join(root, 'md')]
    sys.modules[f'{_PKG}.md'] = md
    _inert_module(f'{_PKG}.md.runner')
    ref_mock = importlib.import_module(f'{_PKG}.md.mock').Mock

    class Mock(ref_mock):
        def run(self, X, tag):
            self.build_input_files(None, None, X)
            return self.read_output()
    md.Mock, md.LennardJones, md.GoldAlkane = Mock, type('LennardJones', (), {}), type('GoldAlkane', (), {})
    db = importlib.import_module(f'{_PKG}.db')
    from scipy.stats import qmc as _qmc
    from .tinygp_stand_in import Random
    # db.py:313 draws an UNSEEDED Latin hypercube / Sobol sequence: here from the generator of the latest jax.random.key
    db.qmc = types.SimpleNamespace(scale=_qmc.scale,
                                   LatinHypercube=lambda d: _qmc.LatinHypercube(d=d, seed=Random.current),
                                   Sobol=lambda d, **k: _qmc.Sobol(d=d, seed=Random.current, **k))
    problem = importlib.import_module(f'{_PKG}.problem')
    _ref = types.SimpleNamespace(Problem=problem.Problem, Database=db.Database, Mock=Mock,
                                 read_yaml_input=sys.modules[f'{_PKG}.io'].read_yaml_input)
    return _ref


def reference_problem():
    return reference_package().Problem


def release():
    """Takes the stand-ins and the reference package out of sys.modules again (the loaded classes keep working: they hold
    what they imported), so that nothing else in the process mistakes a stand-in for an installed package."""
    global _ref
    for name in list(sys.modules):
        if name == _PKG or name.startswith(_PKG + '.') or name in _installed:
            del sys.modules[name]
    _installed.clear()
    _ref = None
