"""Import the reference's NumPy-only modules by file path -- TEST INFRASTRUCTURE ONLY.

``import GaPFlow`` fails here (GaPFlow/__init__.py:25 needs JAX), but
GaPFlow/integrate.py, models/pressure.py, models/sound.py, models/viscous.py and
models/viscosity.py depend only on NumPy/SciPy.  They are loaded straight from
/root/reference, which exists only in the build container: nothing that runs on
the GPU box may call this module.
"""
import importlib.util
import os

REF_ROOT = os.environ.get("GAPFLOW_REFERENCE", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REF_ROOT, "GaPFlow", "integrate.py"))


def load(relpath, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF_ROOT, relpath))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_with_stand_ins(relpath, name, stand_ins):
    """``load`` for a reference module that imports packages which are not installed but which the functions under test
    never call: ``stand_ins`` {module name: module object} are in sys.modules only while the file is executed."""
    import sys
    put = [k for k in stand_ins if k not in sys.modules]
    for k in put:
        sys.modules[k] = stand_ins[k]
    try:
        return load(relpath, name)
    finally:
        for k in put:
            del sys.modules[k]


def modules():
    """Returns dict(integrate, pressure, sound, viscous, viscosity) of reference modules."""
    return dict(
        integrate=load("GaPFlow/integrate.py", "_ref_integrate"),
        pressure=load("GaPFlow/models/pressure.py", "_ref_pressure"),
        sound=load("GaPFlow/models/sound.py", "_ref_sound"),
        viscous=load("GaPFlow/models/viscous.py", "_ref_viscous"),
        viscosity=load("GaPFlow/models/viscosity.py", "_ref_viscosity"),
    )
