"""Load the reference's OWN functions from /root/reference for oracle pinning (build container only).

TEST INFRASTRUCTURE.  Nothing is copied into this repo: ``main_fun`` is imported as a module (matplotlib
stubbed) and the function definitions that live in scripts with import-time side effects
(file I/O, plotting, pysr) are lifted by parsing the script with ``ast`` and exec'ing ONLY the named ``def``
nodes in a scratch namespace.  /root/reference does not exist on the GPU box; tests that need this module
skip there and rely on the fixtures written by make_golden.py.
"""
from __future__ import annotations

import ast
import os
import sys
import types

REF = os.environ.get("CAT_REFERENCE", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REF, "main_fun.py"))


def _stub(name):
    if name not in sys.modules:
        m = types.ModuleType(name)
        m.__dict__.setdefault("__path__", [])
        sys.modules[name] = m
    return sys.modules[name]


def main_fun():
    """The reference's main_fun module (rodrigues_rotation, transform_catenary, solve_catenary,
    integrate_theta_gamma ...), imported unmodified."""
    if "main_fun" in sys.modules and getattr(sys.modules["main_fun"], "__file__", "").startswith(REF):
        return sys.modules["main_fun"]
    mpl = _stub("matplotlib")
    plt = _stub("matplotlib.pyplot")
    mpl.pyplot = plt
    mpl.use = lambda *a, **k: None
    os.environ.setdefault("WANDB_MODE", "disabled")
    os.environ.setdefault("WANDB_SILENT", "true")
    sys.path.insert(0, REF)
    try:
        import importlib
        mod = importlib.import_module("main_fun")
    finally:
        sys.path.remove(REF)
    return mod


def lift(rel_path, names, extra_globals=None):
    """exec the ``def`` nodes called ``names`` from REF/rel_path; returns {name: function}."""
    import numpy as np
    path = os.path.join(REF, rel_path)
    with open(path, encoding="utf-8") as fh:
        tree = ast.parse(fh.read(), filename=path)
    want = set(names)
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in want]
    missing = want - {n.name for n in body}
    if missing:
        raise KeyError(f"{rel_path}: no def {sorted(missing)}")
    ns = {"np": np, "__name__": "reference_lifted"}
    ns.update(extra_globals or {})
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), ns)
    return {n: ns[n] for n in names}


def extract_features18():
    """simple.py:15-41."""
    return lift("simple.py", ["extract_features"])["extract_features"]


def rk4_integration():
    """simulate_rk4_theta_gamma.py:52-68."""
    return lift("simulate_rk4_theta_gamma.py", ["rk4_integration"])["rk4_integration"]


def compute_catenary_3D():
    """models/catenary_3d.py:5-39."""
    return lift(os.path.join("models", "catenary_3d.py"), ["compute_catenary_3D"])["compute_catenary_3D"]


def scaler():
    """saved_models/scaler.pkl (sklearn StandardScaler, simple.py:57-58,114)."""
    import warnings

    import joblib
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return joblib.load(os.path.join(REF, "saved_models", "scaler.pkl"))
