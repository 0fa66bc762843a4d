"""Run-time equation programs: any PySR hall-of-fame row evaluated on the GPU without
rebuilding the library.

The reference evaluates a trained model through ``PySRRegressor.predict`` = a sympy-lambdified ``sympy_format``
string (simple.py:92,105; test_temp.py:53-54), or through ``eval(lambda_format)`` (dd_test_cluster.py:62-69).
``compile_expression`` parses that string (or PySR's own ``Equation`` column: ``^``, ``square``, ``neg`` ...) into
the postfix program of csrc/cat_kernels_expr.cuh; ``ExpressionModel`` has the ``predict`` of the reference's
model objects and plugs into ``rk4_integration`` / ``integrate_theta_gamma`` / ``integrate_second_order``.
"""
from __future__ import annotations

import csv
import ctypes as C
import glob
import os
import re

import numpy as np

from . import _lib
from .context import default_context

OPS = {"const": 0, "var": 1, "+": 2, "-": 3, "*": 4, "/": 5, "neg": 6, "sin": 7, "cos": 8, "tanh": 9, "exp": 10,
       "log": 11, "sqrt": 12, "abs": 13, "square": 14, "cube": 15, "sign": 16, "pow": 17, "powi": 18}
_FUNCS = {"sin": "sin", "cos": "cos", "tanh": "tanh", "exp": "exp", "log": "log", "sqrt": "sqrt", "abs": "abs",
          "Abs": "abs", "square": "square", "cube": "cube", "neg": "neg", "sign": "sign"}
MAX_STACK, MAX_OPS = 32, 2048

_TOKEN = re.compile(r"\s*(?:(\d+\.?\d*(?:[eE][+-]?\d+)?|\.\d+(?:[eE][+-]?\d+)?)|([A-Za-z_]\w*)|(\*\*|\^|[-+*/(),]))")


class ExpressionError(ValueError):
    pass


def _tokens(s):
    out, pos = [], 0
    s = s.strip()
    while pos < len(s):
        m = _TOKEN.match(s, pos)
        if not m:
            raise ExpressionError(f"bad token at {pos}: {s[pos:pos + 16]!r}")
        num, ident, op = m.groups()
        out.append(("num", num) if num is not None else ("id", ident) if ident is not None else ("op", op))
        pos = m.end()
    return out


class _Parser:
    """Recursive descent with Python/sympy precedence (unary minus binds looser than ** / ^).  AST nodes:
    ('num', float) ('var', i) ('neg', a) ('bin', op, a, b) ('call', fn, a) ('powi', a, n) ('pow', a, b)."""

    def __init__(self, toks, names):
        self.t, self.i, self.names = toks, 0, names

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else (None, None)

    def take(self):
        tok = self.peek()
        self.i += 1
        return tok

    def expect(self, op):
        if self.take() != ("op", op):
            raise ExpressionError(f"expected {op!r}")

    def expr(self):
        node = self.term()
        while self.peek() in (("op", "+"), ("op", "-")):
            op = self.take()[1]
            node = ("bin", op, node, self.term())
        return node

    def term(self):
        node = self.unary()
        while self.peek() in (("op", "*"), ("op", "/")):
            op = self.take()[1]
            node = ("bin", op, node, self.unary())
        return node

    def unary(self):
        if self.peek() == ("op", "-"):
            self.take()
            a = self.unary()
            return ("num", -a[1]) if a[0] == "num" else ("neg", a)
        if self.peek() == ("op", "+"):
            self.take()
            return self.unary()
        return self.power()

    def power(self):
        base = self.atom()
        if self.peek() in (("op", "**"), ("op", "^")):
            self.take()
            ex = self.unary()                              # right-associative, exponent may carry a sign
            if ex[0] == "num" and float(ex[1]) == int(float(ex[1])) and abs(int(float(ex[1]))) <= 64:
                return ("powi", base, int(float(ex[1])))
            return ("pow", base, ex)
        return base

    def atom(self):
        kind, val = self.take()
        if kind == "num":
            return ("num", float(val))
        if kind == "id":
            if self.peek() == ("op", "("):
                if val not in _FUNCS:
                    raise ExpressionError(f"unknown function {val!r}")
                self.take()
                a = self.expr()
                self.expect(")")
                return ("call", _FUNCS[val], a)
            if val in self.names:
                return ("var", self.names[val])
            m = re.fullmatch(r"x(\d+)", val)
            if m:
                return ("var", int(m.group(1)))
            raise ExpressionError(f"unknown identifier {val!r}")
        if (kind, val) == ("op", "("):
            a = self.expr()
            self.expect(")")
            return a
        raise ExpressionError(f"unexpected token {val!r}")


def parse(s, feature_names=None):
    names = {n: i for i, n in enumerate(feature_names)} if feature_names else {}
    p = _Parser(_tokens(str(s)), names)
    node = p.expr()
    if p.i != len(p.t):
        raise ExpressionError(f"trailing tokens in {s!r}")
    return node


class Program:
    """Postfix program: ops (n,2) int32 [(opcode, arg)], consts (m,) float64."""

    def __init__(self, ops, consts, n_features, source=""):
        self.ops = np.ascontiguousarray(ops, np.int32).reshape(-1, 2)
        self.consts = np.ascontiguousarray(consts, np.float64)
        self.n_features = int(n_features)
        self.source = source


def compile_expression(s, n_features=None, feature_names=None):
    """Equation string -> Program.  Variables are ``x<i>`` or entries of ``feature_names``."""
    node = parse(s, feature_names)
    ops, consts, cidx = [], [], {}
    used = [-1]

    def const(v):
        k = np.float64(v).tobytes()
        if k not in cidx:
            cidx[k] = len(consts)
            consts.append(float(v))
        return cidx[k]

    def walk(n):
        k = n[0]
        if k == "num":
            ops.append((OPS["const"], const(n[1])))
        elif k == "var":
            used[0] = max(used[0], n[1])
            ops.append((OPS["var"], n[1]))
        elif k == "neg":
            walk(n[1])
            ops.append((OPS["neg"], 0))
        elif k == "bin":
            walk(n[2])
            walk(n[3])
            ops.append((OPS[n[1]], 0))
        elif k == "call":
            walk(n[2])
            ops.append((OPS[n[1]], 0))
        elif k == "powi":
            walk(n[1])
            ops.append((OPS["powi"], n[2]))
        elif k == "pow":
            walk(n[1])
            walk(n[2])
            ops.append((OPS["pow"], 0))
        else:
            raise ExpressionError(k)

    walk(node)
    depth = peak = 0
    for op, _ in ops:
        depth += 1 if op in (0, 1) else -1 if op in (2, 3, 4, 5, 17) else 0
        peak = max(peak, depth)
    if peak > MAX_STACK or len(ops) > MAX_OPS:
        raise ExpressionError(f"expression too large (stack {peak} > {MAX_STACK} or {len(ops)} ops > {MAX_OPS})")
    nf = used[0] + 1 if n_features is None else int(n_features)
    if feature_names and n_features is None:
        nf = len(feature_names)
    if used[0] >= nf:
        raise ExpressionError(f"expression reads feature {used[0]} but n_features = {nf}")
    return Program(ops, consts, max(nf, 1), str(s))


class ExpressionModel:
    """``predict(X)`` of the reference's model objects (PySRRegressor / lambda_format callables) for ANY equation
    string; optional StandardScaler constants are applied first (test_temp.py:49)."""

    def __init__(self, expression, n_features=None, feature_names=None, scaler_mean=None, scaler_scale=None,
                 ctx=None):
        self.program = expression if isinstance(expression, Program) else compile_expression(
            expression, n_features, feature_names)
        self.mean = None if scaler_mean is None else np.ascontiguousarray(scaler_mean, np.float64)
        self.scale = None if scaler_scale is None else np.ascontiguousarray(scaler_scale, np.float64)
        if (self.mean is None) != (self.scale is None):
            raise ValueError("scaler_mean and scaler_scale go together")
        if self.mean is not None and (self.mean.shape != (self.program.n_features,) or
                                      self.scale.shape != (self.program.n_features,)):
            raise ValueError("scaler constants must have n_features entries")
        self._ctx = ctx

    @property
    def ctx(self):
        return self._ctx or default_context()

    def predict(self, X, dtype=np.float64):
        dtype = np.dtype(dtype)
        X = np.asarray(getattr(X, "values", X), dtype)
        if X.ndim == 1:
            X = X[None]
        pr = self.program
        if X.ndim != 2 or X.shape[1] != pr.n_features:
            raise ValueError(f"X must be (n,{pr.n_features})")
        X = np.ascontiguousarray(X)
        out = np.empty(X.shape[0], dtype)
        fn = self.ctx.lib.cat_expr_eval_f64 if dtype == np.float64 else self.ctx.lib.cat_expr_eval_f32
        self.ctx.check(fn(self.ctx.handle, X.shape[0], pr.n_features, _lib.ptr(X), len(pr.ops), _lib.ptr(pr.ops),
                          len(pr.consts), _lib.ptr(pr.consts) if len(pr.consts) else None, _lib.ptr(self.mean),
                          _lib.ptr(self.scale), _lib.ptr(out)))
        return out

    __call__ = predict

    def integrate(self, X, time, y0, method="rk4", return_rate=False):
        """Open-loop integration of dy/dt = f(X[i]) over one trajectory (T,F) or a batch (n,T,F).
        method 'rk4' (simulate_rk4_theta_gamma.py:52-68), 'euler' (main_fun.py:735-764) or 'second_order'
        (the equation is the SECOND derivative; test_cluster.py:110-129)."""
        m = {"rk4": 0, "euler": 1, "second_order": 2}[method]
        X = np.asarray(getattr(X, "values", X), np.float64)
        single = X.ndim == 2
        if single:
            X = X[None]
        pr = self.program
        if X.ndim != 3 or X.shape[2] != pr.n_features:
            raise ValueError(f"X must be (T,{pr.n_features}) or (n,T,{pr.n_features})")
        if self.mean is not None:                        # the model's StandardScaler applies to integration too
            X = (X - self.mean) / self.scale
        X = np.ascontiguousarray(X)
        n, T = X.shape[0], X.shape[1]
        time = np.ascontiguousarray(time, np.float64)
        if time.shape != (T,):
            raise ValueError("time must be (T,)")
        y0 = np.ascontiguousarray(np.broadcast_to(np.asarray(y0, np.float64), (n,)))
        y = np.empty((n, T))
        yd = np.empty((n, T)) if (m == 2 and return_rate) else None
        self.ctx.check(self.ctx.lib.cat_expr_integrate_f64(
            self.ctx.handle, n, T, pr.n_features, _lib.ptr(X), _lib.ptr(time), _lib.ptr(y0), len(pr.ops),
            _lib.ptr(pr.ops), len(pr.consts), _lib.ptr(pr.consts) if len(pr.consts) else None, m, _lib.ptr(y),
            _lib.ptr(yd)))
        y = y[0] if single else y
        if return_rate:
            return y, (yd[0] if (single and yd is not None) else yd)
        return y


def integrate_second_order(model, X, time, y0):
    """theta from a model of its SECOND derivative: rate starts at 0 and is Euler-integrated, then the angle
    from y0 (test_cluster.py:110-129, dd_test_cluster.py:80-99)."""
    return model.integrate(X, time, y0, method="second_order")


def load_equation_table(path):
    """Rows of a PySR equation file: the reference's ``d*_results_*.csv`` / ``equations_*.csv``
    (complexity, loss, score, equation, sympy_format, lambda_format; simple.py:92,105) or PySR's own
    ``hall_of_fame_*.csv`` (Complexity, Loss, Equation).  -> list of dict(complexity, loss, expression)."""
    rows = []
    with open(path, newline="") as fh:
        for r in csv.DictReader(fh):
            expr = r.get("sympy_format") or r.get("Equation") or r.get("equation")
            if not expr:
                continue
            cx = r.get("complexity") or r.get("Complexity")
            loss = r.get("loss") or r.get("Loss")
            rows.append(dict(complexity=int(float(cx)) if cx else None, loss=float(loss) if loss else None,
                             expression=expr))
    return rows


def load_run(run_dir, feature_names=None, n_features=None):
    """Every equation table of one training run directory (cluster_run/outputs/<run>/ layout) ->
    {file name: [(complexity, loss, ExpressionModel)]}."""
    out = {}
    pats = ("*results*.csv", "equations*.csv", os.path.join("*", "hall_of_fame*.csv"))
    for pat in pats:
        for path in sorted(glob.glob(os.path.join(run_dir, pat))):
            models = []
            for r in load_equation_table(path):
                models.append((r["complexity"], r["loss"],
                               ExpressionModel(r["expression"], n_features, feature_names)))
            out[os.path.relpath(path, run_dir)] = models
    return out
