"""Importable alias of the package directory ``tensorflow-distributed-clustering_b200`` (its
name has a hyphen, so a plain ``import`` statement cannot name it)."""
import importlib
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_pkg = importlib.import_module("tensorflow-distributed-clustering_b200")
globals().update({k: getattr(_pkg, k) for k in _pkg.__all__})
_lib = _pkg._lib
package = _pkg
__all__ = list(_pkg.__all__) + ["package"]
