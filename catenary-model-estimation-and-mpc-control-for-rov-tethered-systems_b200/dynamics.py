"""Reference call surface for the theta/gamma dynamics half of the path, backed by the CUDA library.

  extract_features(df)                                  simple.py:15-41  (18 columns)
  EquationModel(name).predict(X)                        PySRRegressor.predict stand-in (test_temp.py:53-54)
  rk4_integration(model, x_input, time, y0)             simulate_rk4_theta_gamma.py:52-68
  integrate_theta_gamma(mt, mg, X, time, th0, ga0)      main_fun.py:735-764
"""
from __future__ import annotations

import ctypes as C
import json
import os

import numpy as np

from . import _lib
from .context import default_context

_HERE = os.path.dirname(os.path.abspath(__file__))
_MODEL_JSON = os.path.join(_HERE, "artifacts", "model.json")
_model = None


def load_model():
    global _model
    if _model is None:
        with open(_MODEL_JSON) as fh:
            _model = json.load(fh)
    return _model


def equation_table(ctx=None):
    """[(eq_id, 'dtheta_dt'|'dgamma_dt', complexity, live feature indices)] as compiled into the library."""
    lib = (ctx or default_context()).lib if ctx else _lib.load()
    out = []
    for i in range(lib.cat_equation_count()):
        g, cx, mask = C.c_int(), C.c_int(), C.c_uint32()
        lib.cat_equation_info(i, C.byref(g), C.byref(cx), C.byref(mask))
        out.append((i, "dgamma_dt" if g.value else "dtheta_dt", cx.value,
                    [j for j in range(18) if mask.value >> j & 1]))
    return out


def equation_id(name, complexity=None):
    """Generated equation id of a hall-of-fame row (None = the shipped row of eq_*.txt)."""
    lib = _lib.load()
    if name not in ("dtheta_dt", "dgamma_dt"):
        raise ValueError("name must be 'dtheta_dt' or 'dgamma_dt'")
    if complexity is None:
        return lib.cat_equation_shipped(0 if name == "dtheta_dt" else 1)
    for i, nm, cx, _ in equation_table():
        if nm == name and cx == complexity:
            return i
    raise ValueError(f"no {name} equation of complexity {complexity}")


def extract_features(df, *, scaled=False, ctx=None):
    """simple.py:15-41 on a DataFrame / mapping with the reference's column names (data units mm, mm/s)."""
    ctx = ctx or default_context()
    g = lambda k: np.asarray(df[k], np.float64)
    P0 = np.ascontiguousarray(np.stack([g("rod_end X"), g("rod_end Y"), g("rod_end Z")], axis=1) / 1000)
    P1 = np.ascontiguousarray(np.stack([g("robot_cable_attach_point X"), g("robot_cable_attach_point Y"),
                                        g("robot_cable_attach_point Z")], axis=1) / 1000)
    V1 = np.ascontiguousarray(np.stack([g("rob_cor_speed X"), g("rob_cor_speed Y"), g("rob_cor_speed Z")], axis=1))
    t = np.ascontiguousarray(g("Time"))
    th = np.ascontiguousarray(g("Theta"))
    ga = np.ascontiguousarray(g("Gamma"))
    T = t.shape[0]
    X = np.empty((T, 18))
    ctx.check(ctx.lib.cat_features_f64(ctx.handle, T, _lib.ptr(P0), _lib.ptr(P1), _lib.ptr(V1), _lib.ptr(t),
                                       _lib.ptr(th), _lib.ptr(ga), int(bool(scaled)), _lib.ptr(X)))
    return X


def predict(X, eq_id, *, apply_scaler=False, dtype=np.float64, ctx=None):
    ctx = ctx or default_context()
    dtype = np.dtype(dtype)
    X = np.ascontiguousarray(X, dtype)
    if X.shape[-1] != 18:
        raise ValueError("X must have 18 columns")
    lead = X.shape[:-1]
    X2 = X.reshape(-1, 18)
    out = np.empty(X2.shape[0], dtype)
    fn = ctx.lib.cat_predict_f64 if dtype == np.float64 else ctx.lib.cat_predict_f32
    ctx.check(fn(ctx.handle, X2.shape[0], _lib.ptr(X2), int(eq_id), int(bool(apply_scaler)), _lib.ptr(out)))
    return out.reshape(lead)


class EquationModel:
    """Object with the ``predict`` surface the reference's integrators call on a PySRRegressor."""

    def __init__(self, name, complexity=None, ctx=None):
        self.name = name
        self.eq_id = equation_id(name, complexity)
        self._ctx = ctx

    def predict(self, X):
        return predict(np.asarray(X, np.float64), self.eq_id, ctx=self._ctx)


def _integrate(model, x_input, time, y0, method, dtype, ctx):
    if hasattr(model, "program") and hasattr(model, "integrate"):      # expr.ExpressionModel: run-time program
        if np.dtype(dtype) != np.float64:
            raise ValueError("run-time equation programs integrate in float64")
        return model.integrate(x_input, time, y0, method="rk4" if method == 0 else "euler")
    ctx = ctx or default_context()
    if not isinstance(model, EquationModel):
        raise TypeError("model must be an EquationModel (generated at build time) or an expr.ExpressionModel")
    dtype = np.dtype(dtype)
    X = np.ascontiguousarray(x_input, dtype)
    single = X.ndim == 2
    X3 = X.reshape((-1,) + X.shape[-2:])
    n, T, F = X3.shape
    if F != 18:
        raise ValueError("x_input must have 18 columns")
    t = np.ascontiguousarray(time, dtype)
    if t.shape != (T,):
        raise ValueError("time must have one entry per row of x_input")
    y0a = np.ascontiguousarray(np.broadcast_to(np.asarray(y0, dtype), (n,)))
    y = np.empty((n, T), dtype)
    fn = ctx.lib.cat_rk4_open_f64 if dtype == np.float64 else ctx.lib.cat_rk4_open_f32
    ctx.check(fn(ctx.handle, n, T, _lib.ptr(X3), _lib.ptr(t), _lib.ptr(y0a), model.eq_id, method, _lib.ptr(y)))
    return y[0] if single else y


def rk4_integration(model, x_input, time, y0, *, dtype=np.float64, ctx=None):
    """simulate_rk4_theta_gamma.py:52-68.  x_input (T,18) -> (T,), or (n,T,18) -> (n,T) for n trajectories."""
    return _integrate(model, x_input, time, y0, 0, dtype, ctx)


def integrate_theta_gamma(model_theta, model_gamma, X, time_array, theta_0, gamma_0, *, ctx=None):
    """main_fun.py:735-764 (explicit Euler, open loop) -> (theta, gamma)."""
    return (_integrate(model_theta, X, time_array, theta_0, 1, np.float64, ctx),
            _integrate(model_gamma, X, time_array, gamma_0, 1, np.float64, ctx))
