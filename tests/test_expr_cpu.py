"""Host side of the run-time equation evaluator (expr.py): every equation string of the reference's PySR
outputs (tests/golden/equation_corpus.*, made by oracle/make_equation_corpus.py with sympy.lambdify / Python's
own parser) compiles to a well-formed postfix program whose NumPy interpretation reproduces the golden values."""
import importlib
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN

PKG = "catenary-model-estimation-and-mpc-control-for-rov-tethered-systems_b200"


def corpus():
    with open(os.path.join(GOLDEN, "equation_corpus.json")) as fh:
        meta = json.load(fh)
    y = np.load(os.path.join(GOLDEN, "equation_corpus.npz"))["y"]
    return meta, y


def corpus_inputs(meta, i):
    it = meta["items"][i]
    return np.random.default_rng(meta["seed_base"] + i).normal(size=(meta["n_rows"], it["n_features"])) * meta["scale"]


def run_program(prog, X):
    """NumPy interpretation of a postfix program (the checker of the COMPILE step; the GPU runs the same ops)."""
    st = []
    with np.errstate(all="ignore"):
        for op, arg in prog.ops:
            if op == 0:
                st.append(np.full(X.shape[0], prog.consts[arg]))
            elif op == 1:
                st.append(X[:, arg].copy())
            elif op in (2, 3, 4, 5, 17):
                b = st.pop()
                a = st.pop()
                st.append(a + b if op == 2 else a - b if op == 3 else a * b if op == 4 else a / b if op == 5
                          else np.power(a, b))
            else:
                a = st.pop()
                st.append({6: np.negative, 7: np.sin, 8: np.cos, 9: np.tanh, 10: np.exp, 11: np.log, 12: np.sqrt,
                           13: np.abs, 14: np.square, 15: lambda v: v * v * v, 16: np.sign,
                           18: lambda v: np.power(v, float(arg))}[op](a))
    assert len(st) == 1
    return st[0]


def close(got, want, rtol=1e-9):
    if not np.array_equal(np.isnan(got), np.isnan(want)):
        return False
    m = np.isfinite(want)
    if not np.array_equal(got[~m & ~np.isnan(want)], want[~m & ~np.isnan(want)]):
        return False
    return bool(np.all(np.abs(got[m] - want[m]) <= rtol * np.abs(want[m]) + 1e-300))


def test_every_reference_equation_compiles_and_matches():
    ex = importlib.import_module(PKG + ".expr")
    meta, y = corpus()
    assert len(meta["items"]) == y.shape[0] >= 3000
    bad = []
    for i, it in enumerate(meta["items"]):
        prog = ex.compile_expression(it["expr"], it["n_features"], it["feature_names"])
        got = run_program(prog, corpus_inputs(meta, i))
        if not close(got, y[i], 1e-9):
            bad.append((i, it["expr"], got, y[i]))
    assert not bad, f"{len(bad)} of {len(meta['items'])} differ, first: {bad[0]}"


def test_parser_precedence_and_errors():
    ex = importlib.import_module(PKG + ".expr")
    X = np.array([[0.7, -1.3, 2.0]])

    def val(s, **kw):
        return float(run_program(ex.compile_expression(s, 3, **kw), X)[0])

    assert val("-x0**2") == -(0.7 ** 2)
    assert val("-x0^2") == -(0.7 ** 2)
    assert val("2**-2") == 0.25
    assert val("x2**-1") == 0.5
    assert val("x0 - x1 - x2") == (0.7 - -1.3) - 2.0
    assert val("x0 / x1 / x2") == (0.7 / -1.3) / 2.0
    assert val("neg(square(x1)) + Abs(x1)") == -(1.3 ** 2) + 1.3
    assert val("a*b + c", feature_names=["a", "b", "c"]) == 0.7 * -1.3 + 2.0
    assert val("1.5e-3 * x0") == 1.5e-3 * 0.7
    assert abs(val("x2**0.5") - 2.0 ** 0.5) < 1e-15
    for s in ("x0 +", "foo(x0)", "x0 $ x1", "(x0", "x7", "y"):
        with pytest.raises(ex.ExpressionError):
            ex.compile_expression(s, 3)
    deep = "(" * 40 + "x0" + ")" * 40
    assert val(deep) == 0.7                                  # parentheses alone do not deepen the stack
    chain = "x0" + "".join(f"+(x1*(x2+{i}" for i in range(40)) + "))" * 40
    with pytest.raises(ex.ExpressionError):
        ex.compile_expression(chain, 3)                      # stack deeper than the device's 32 slots


def test_equation_table_loaders(tmp_path):
    ex = importlib.import_module(PKG + ".expr")
    a = tmp_path / "dtheta_results_1.csv"
    a.write_text("complexity,loss,score,equation,sympy_format,lambda_format\n"
                 "3,0.1,0.0,(x4 * 0.5),0.5*x4,PySRFunction(X=>0.5*x4)\n"
                 "5,0.05,0.1,((x4 * 0.5) - x1),0.5*x4 - x1,PySRFunction(X=>0.5*x4 - x1)\n")
    d = tmp_path / "dgamma_dt"
    d.mkdir()
    (d / "hall_of_fame_2025.csv").write_text("Complexity,Loss,Equation\n1,0.2,\"1.5e-8\"\n4,0.1,\"square(x2) ^ 2\"\n")
    rows = ex.load_equation_table(str(a))
    assert [r["complexity"] for r in rows] == [3, 5] and rows[1]["expression"] == "0.5*x4 - x1"
    run = ex.load_run(str(tmp_path), n_features=6)
    assert sorted(run) == ["dgamma_dt/hall_of_fame_2025.csv", "dtheta_results_1.csv"]
    assert [m[0] for m in run["dgamma_dt/hall_of_fame_2025.csv"]] == [1, 4]
