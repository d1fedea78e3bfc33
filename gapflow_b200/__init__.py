"""B200-native hot path of GaPFlow: FV update and GP inference behind the reference's Python API.

Names resolve lazily so that ``import gapflow_b200`` stays cheap (``gapflow_b200.build`` must be
importable before the shared library exists).
"""
_LAZY = {
    "Problem": ("gapflow_b200.problem", "Problem"),            # GaPFlow/problem.py:41
    "FVSolver": ("gapflow_b200.fv", "FVSolver"),
    "GPModel": ("gapflow_b200.gp", "GPModel"),                 # GaPFlow/models/gp.py:41
    "SlabSolver": ("gapflow_b200.dist", "SlabSolver"),
}

__all__ = sorted(_LAZY)


def __getattr__(name):
    if name in _LAZY:
        import importlib
        mod, attr = _LAZY[name]
        return getattr(importlib.import_module(mod), attr)
    raise AttributeError(f"module 'gapflow_b200' has no attribute {name!r}")
