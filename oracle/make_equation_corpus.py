#!/usr/bin/env python
"""Golden corpus for the run-time equation evaluator (TEST INFRASTRUCTURE; needs /root/reference, run here only).

Collects every equation row of the reference's PySR outputs -- ``*results*.csv`` / ``equations*.csv``
(``sympy_format`` column, what PySRRegressor.predict lambdifies; simple.py:92,105) and ``hall_of_fame*.csv``
(PySR's own ``Equation`` column) -- de-duplicates them, and evaluates each on seeded random feature rows:
  * sympy_format rows: sympy.sympify + lambdify(modules='numpy'), i.e. the arithmetic PySR's predict runs;
  * Equation rows (Julia operator names): Python's own parser via eval() over a NumPy namespace.
Writes tests/golden/equation_corpus.json (strings, variable names) and equation_corpus.npz (inputs seeds are
implicit: X = default_rng(1000 + i).normal(size=(8, F)) * 0.8; outputs (n_eq, 8))."""
import csv
import glob
import json
import os
import re
import sys

import numpy as np
import sympy

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
N_ROWS = 8
NS = {"sin": np.sin, "cos": np.cos, "tanh": np.tanh, "exp": np.exp, "log": np.log, "sqrt": np.sqrt, "abs": np.abs,
      "Abs": np.abs, "square": np.square, "cube": lambda v: v * v * v, "neg": np.negative, "sign": np.sign}
SYM = {"square": lambda v: v ** 2, "cube": lambda v: v ** 3, "neg": lambda v: -v}


def variables(expr):
    names = set(re.findall(r"[A-Za-z_]\w*", re.sub(r"\d+\.?\d*[eE][+-]?\d+", "0", expr)))
    return sorted(n for n in names if n not in NS)


def main():
    seen, items = set(), []
    files = sorted(glob.glob(os.path.join(REF, "**", "*results*.csv"), recursive=True) +
                   glob.glob(os.path.join(REF, "**", "equations*.csv"), recursive=True) +
                   glob.glob(os.path.join(REF, "**", "hall_of_fame*.csv"), recursive=True))
    for f in files:
        try:
            rows = list(csv.DictReader(open(f, newline="")))
        except Exception:
            continue
        for r in rows:
            for col, kind in (("sympy_format", "sympy"), ("Equation", "pysr"), ("equation", "pysr")):
                s = r.get(col)
                if not s or (kind, s) in seen:
                    continue
                seen.add((kind, s))
                items.append(dict(kind=kind, expr=s, file=os.path.relpath(f, REF)))
    out_items, outs = [], []
    skipped = 0
    for it in items:
        names = variables(it["expr"])
        xs = [n for n in names if re.fullmatch(r"x\d+", n)]
        if xs and len(xs) == len(names):
            F = max(int(n[1:]) for n in xs) + 1
            cols = {f"x{j}": j for j in range(F)}
            fnames = None
        else:
            fnames = names
            F = max(len(names), 1)
            cols = {n: j for j, n in enumerate(names)}
        X = np.random.default_rng(1000 + len(out_items)).normal(size=(N_ROWS, F)) * 0.8
        env = dict(NS)
        env.update({n: X[:, j] for n, j in cols.items()})
        try:
            with np.errstate(all="ignore"):
                if it["kind"] == "sympy":
                    syms = {n: sympy.Symbol(n) for n in cols}
                    e = sympy.sympify(it["expr"], locals={**syms, **SYM})
                    order = sorted(cols, key=cols.get)
                    fn = sympy.lambdify([syms[n] for n in order], e, modules="numpy")
                    y = np.broadcast_to(np.asarray(fn(*[X[:, cols[n]] for n in order]), np.float64), (N_ROWS,))
                else:
                    y = np.broadcast_to(np.asarray(eval(it["expr"].replace("^", "**"), {"__builtins__": {}}, env),
                                                   np.float64), (N_ROWS,))
        except Exception as ex:                              # a row neither evaluator can read is not golden
            skipped += 1
            continue
        out_items.append(dict(kind=it["kind"], expr=it["expr"], file=it["file"], n_features=F,
                              feature_names=fnames))
        outs.append(np.array(y, np.float64))
    with open(os.path.join(OUT, "equation_corpus.json"), "w") as fh:
        json.dump(dict(n_rows=N_ROWS, seed_base=1000, scale=0.8, items=out_items), fh)
    np.savez_compressed(os.path.join(OUT, "equation_corpus.npz"), y=np.stack(outs))
    print(f"{len(out_items)} equations from {len(files)} files ({skipped} unreadable rows skipped)")


if __name__ == "__main__":
    main()
