"""Importable alias of the package directory ``editable-gaussian-reflections_b200`` (whose name is not a Python
identifier): ``import egr_b200`` == that package."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("editable-gaussian-reflections_b200")
sys.modules[__name__] = _pkg
