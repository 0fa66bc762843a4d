"""Short import alias: ``import catenary_b200`` == the package in
``catenary-model-estimation-and-mpc-control-for-rov-tethered-systems_b200/`` (a directory name that is not a
Python identifier)."""
import importlib
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_pkg = importlib.import_module("catenary-model-estimation-and-mpc-control-for-rov-tethered-systems_b200")
sys.modules[__name__] = _pkg
