"""Build-time code generation: PySR hall-of-fame equations + scaler -> inlined __device__ functions.

Input : artifacts/model.json (written by tools/extract_artifacts.py from the reference's
        saved_models/equations_dtheta_dt.csv, equations_dgamma_dt.csv, eq_*.txt, scaler.pkl).
Output: csrc/gen/dyn_eqs.cuh

The string that is compiled is the CSV ``sympy_format`` column -- the same string
PySRRegressor.predict lambdifies (reference simple.py:92,105 wrote it), so the association
of the generated code follows what the reference evaluates.  Operator vocabulary:
+ - * / ** (integer powers), sin cos tanh exp log sqrt Abs (simple.py:65-66 plus the other runs).
"""
from __future__ import annotations

import json
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
MODEL_JSON = os.path.join(_HERE, "artifacts", "model.json")
OUT_DIR = os.path.join(_HERE, "csrc", "gen")
OUT_FILE = os.path.join(OUT_DIR, "dyn_eqs.cuh")
ICDF_JSON = os.path.join(_HERE, "artifacts", "icdf_table.json")
ICDF_FILE = os.path.join(OUT_DIR, "icdf_table.cuh")

_TOKEN = re.compile(r"\s*(?:(\d+\.?\d*(?:[eE][+-]?\d+)?|\.\d+(?:[eE][+-]?\d+)?)|([A-Za-z_]\w*)|(\*\*|[-+*/()]))")
_FUNCS = {"sin": "sin", "cos": "cos", "tanh": "tanh", "exp": "exp", "log": "log",
          "sqrt": "sqrt", "Abs": "abs", "abs": "abs"}


class ParseError(ValueError):
    pass


def tokenize(s):
    out, pos = [], 0
    s = s.strip()
    while pos < len(s):
        m = _TOKEN.match(s, pos)
        if not m:
            raise ParseError(f"bad token at {pos}: {s[pos:pos + 12]!r}")
        num, ident, op = m.groups()
        out.append(("num", num) if num is not None else ("id", ident) if ident is not None else ("op", op))
        pos = m.end()
    return out


class _Parser:
    """Recursive descent -> nested tuples: ('num',str) ('var',i) ('neg',a) ('bin',op,a,b)
    ('call',fn,a) ('pow',a,n)."""

    def __init__(self, toks):
        self.t, self.i = toks, 0

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else (None, None)

    def take(self):
        tok = self.peek()
        self.i += 1
        return tok

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
            return ("neg", self.unary())
        if self.peek() == ("op", "+"):
            self.take()
            return self.unary()
        return self.power()

    def power(self):
        base = self.atom()
        if self.peek() == ("op", "**"):
            self.take()
            ex = self.unary()
            if ex[0] == "num" and float(ex[1]) == int(float(ex[1])) and 0 < int(float(ex[1])) <= 8:
                return ("pow", base, int(float(ex[1])))
            raise ParseError("only small positive integer powers are supported")
        return base

    def atom(self):
        kind, val = self.take()
        if kind == "num":
            return ("num", val)
        if kind == "id":
            m = re.fullmatch(r"x(\d+)", val)
            if m:
                return ("var", int(m.group(1)))
            if val in _FUNCS:
                if self.take() != ("op", "("):
                    raise ParseError(f"expected ( after {val}")
                a = self.expr()
                if self.take() != ("op", ")"):
                    raise ParseError("expected )")
                return ("call", _FUNCS[val], a)
            raise ParseError(f"unknown identifier {val}")
        if (kind, val) == ("op", "("):
            a = self.expr()
            if self.take() != ("op", ")"):
                raise ParseError("expected )")
            return a
        raise ParseError(f"unexpected token {val!r}")


def parse(s):
    p = _Parser(tokenize(s))
    node = p.expr()
    if p.i != len(p.t):
        raise ParseError(f"trailing tokens in {s!r}")
    return node


def live_features(node, acc=None):
    acc = set() if acc is None else acc
    if node[0] == "var":
        acc.add(node[1])
    elif node[0] in ("neg",):
        live_features(node[1], acc)
    elif node[0] == "bin":
        live_features(node[2], acc)
        live_features(node[3], acc)
    elif node[0] == "call":
        live_features(node[2], acc)
    elif node[0] == "pow":
        live_features(node[1], acc)
    return acc


def emit(node):
    """C++ expression over ``x[i]`` in type T with math routed through M (CatMath<T>)."""
    k = node[0]
    if k == "num":
        v = repr(float(node[1]))
        return f"T({v})"
    if k == "var":
        return f"x[{node[1]}]"
    if k == "neg":
        return f"(-{emit(node[1])})"
    if k == "bin":
        return f"({emit(node[2])} {node[1]} {emit(node[3])})"
    if k == "call":
        return f"M::{node[1]}_({emit(node[2])})"
    if k == "pow":
        return f"M::template ipow<{node[2]}>({emit(node[1])})"
    raise ParseError(k)


_VEC_FUNCS = {"sin"}     # functions with an interleaved N-wide implementation in CatMath (sin_v)


def _has_call(node):
    k = node[0]
    if k == "call":
        return True
    if k in ("neg", "pow"):
        return _has_call(node[1])
    if k == "bin":
        return _has_call(node[2]) or _has_call(node[3])
    return False


def _leaf_vec_calls(node, acc):
    """Calls of a vectorised function whose argument contains no further call, in evaluation order."""
    k = node[0]
    if k == "call":
        if node[1] in _VEC_FUNCS and not _has_call(node[2]):
            acc.append(node)
        else:
            _leaf_vec_calls(node[2], acc)
    elif k in ("neg", "pow"):
        _leaf_vec_calls(node[1], acc)
    elif k == "bin":
        _leaf_vec_calls(node[2], acc)
        _leaf_vec_calls(node[3], acc)
    return acc


def emit_vec(node, stmts, counter, batched=None):
    """Expression for row ``n`` of an N-row batch.  Transcendental calls are hoisted into statements that
    evaluate all N rows at once (``M::template sin_v<..>``); the innermost calls of the whole expression go
    into ONE batch of M*N arguments, so every polynomial coefficient is fetched once and the M*N Horner
    chains interleave."""
    if batched is None:
        batched = {}
        leaves = _leaf_vec_calls(node, [])
        by_fn = {}
        for lf in leaves:
            by_fn.setdefault(lf[1], []).append(lf)
        for fn, group in by_fn.items():
            name = f"b{counter[0]}"
            counter[0] += 1
            m = len(group)
            stmts.append(f"  T {name}[{m} * N];")
            stmts.append("#pragma unroll")
            stmts.append("  for (int n = 0; n < N; ++n) {")
            for slot, lf in enumerate(group):
                stmts.append(f"    {name}[{slot} * N + n] = {emit_vec(lf[2], stmts, counter, batched)};")
                batched[id(lf)] = f"{name}[{slot} * N + n]"
            stmts.append("  }")
            stmts.append(f"  M::template {fn}_v<{m} * N>({name});")
    k = node[0]
    if k == "num":
        return f"T({float(node[1])!r})"
    if k == "var":
        return f"x[n][{node[1]}]"
    if k == "neg":
        return f"(-{emit_vec(node[1], stmts, counter, batched)})"
    if k == "bin":
        return (f"({emit_vec(node[2], stmts, counter, batched)} {node[1]} "
                f"{emit_vec(node[3], stmts, counter, batched)})")
    if k == "pow":
        return f"M::template ipow<{node[2]}>({emit_vec(node[1], stmts, counter, batched)})"
    if k == "call":
        if id(node) in batched:
            return batched[id(node)]
        arg = emit_vec(node[2], stmts, counter, batched)
        name = f"t{counter[0]}"
        counter[0] += 1
        stmts.append(f"  T {name}[N];")
        stmts.append("#pragma unroll")
        stmts.append(f"  for (int n = 0; n < N; ++n) {name}[n] = {arg};")
        if node[1] in _VEC_FUNCS:
            stmts.append(f"  M::template {node[1]}_v<N>({name});")
        else:
            stmts.append("#pragma unroll")
            stmts.append(f"  for (int n = 0; n < N; ++n) {name}[n] = M::{node[1]}_({name}[n]);")
        return f"{name}[n]"
    raise ParseError(k)


def evaluate(node, x, mod):
    """Host-side evaluation of the AST (used only by codegen self-tests; ``mod`` = math or numpy)."""
    k = node[0]
    if k == "num":
        return float(node[1])
    if k == "var":
        return x[node[1]]
    if k == "neg":
        return -evaluate(node[1], x, mod)
    if k == "bin":
        a, b = evaluate(node[2], x, mod), evaluate(node[3], x, mod)
        return a + b if node[1] == "+" else a - b if node[1] == "-" else a * b if node[1] == "*" else a / b
    if k == "call":
        fn = {"abs": abs}.get(node[1]) or getattr(mod, node[1])
        return fn(evaluate(node[2], x, mod))
    if k == "pow":
        return evaluate(node[1], x, mod) ** node[2]
    raise ParseError(k)


def _mask(feats):
    m = 0
    for i in feats:
        m |= 1 << i
    return m


def generate(model_json=MODEL_JSON, out_file=OUT_FILE):
    with open(model_json) as fh:
        model = json.load(fh)
    nf = model["n_features"]
    mean = model["scaler_mean"]
    scale = model["scaler_scale"]
    lines = [
        "// GENERATED by codegen.py from artifacts/model.json -- do not edit.",
        "// Source data: reference saved_models/equations_dtheta_dt.csv, equations_dgamma_dt.csv,",
        "// eq_dtheta_dt.txt, eq_dgamma_dt.txt (shipped complexity), scaler.pkl (simple.py:57-58,114).",
        "#pragma once",
        f"#define CAT_N_FEATURES {nf}",
        "",
        "// StandardScaler constants: xs = (x - mean) * inv_scale   (test_temp.py:49)",
        "template <typename T> struct CatScaler {",
        f"  static __host__ __device__ constexpr T mean(int i) {{ constexpr double v[{nf}] = {{"
        + ", ".join(repr(float(v)) for v in mean) + "}; return T(v[i]); }",
        f"  static __host__ __device__ constexpr T inv_scale(int i) {{ constexpr double v[{nf}] = {{"
        + ", ".join(repr(1.0 / float(v)) for v in scale) + "}; return T(v[i]); }",
        "  // the same map as one FMA: xs = x * inv_scale + offset, offset = -mean * inv_scale (the rollout kernel)",
        "  static __host__ __device__ constexpr T offset(int i) { return T(-double(CatScaler<double>::mean(i)) * "
        "double(CatScaler<double>::inv_scale(i))); }",
        "};",
        "// (inv_scale, offset) pairs in constant memory: one 128-bit uniform load feeds the FMA of a feature,",
        "// where two 64-bit literals would cost four UMOVs at every use",
        f"__constant__ __align__(16) double CAT_SCALER_FMA[{2 * nf}] = {{"
        + ", ".join(repr(1.0 / float(sc)) + ", " + repr(-float(m) * (1.0 / float(sc))) for m, sc in zip(mean, scale))
        + "};",
        "",
    ]
    table = []          # (eq_id, name, complexity, live mask)
    eq_id = 0
    shipped = {}
    for name in ("dtheta_dt", "dgamma_dt"):
        entry = model[name]
        for row in entry["rows"]:
            node = parse(row["sympy_format"])
            live = sorted(live_features(node))
            if any(i >= nf for i in live):
                raise ParseError(f"{name} cx{row['complexity']} uses feature >= {nf}")
            fn = f"cat_eq_{name}_c{row['complexity']}"
            lines += [
                f"// {name} complexity {row['complexity']} loss {row['loss']}: {row['sympy_format']}",
                f"// live features: {live}",
                "template <typename T, typename M>",
                f"__device__ __forceinline__ T {fn}(const T (&x)[CAT_N_FEATURES]) {{",
                f"  return {emit(node)};",
                "}",
                "",
            ]
            stmts, counter = [], [0]
            expr = emit_vec(node, stmts, counter)
            lines += [
                "template <typename T, typename M, int N>",
                f"__device__ __forceinline__ void {fn}_v(const T (&x)[N][CAT_N_FEATURES], T (&out)[N]) {{",
                *stmts,
                "#pragma unroll",
                f"  for (int n = 0; n < N; ++n) out[n] = {expr};",
                "}",
                "",
            ]
            table.append((eq_id, name, row["complexity"], _mask(live), fn))
            if row["complexity"] == entry["shipped_complexity"]:
                shipped[name] = (eq_id, fn, _mask(live))
            eq_id += 1
    lines += [f"#define CAT_N_EQUATIONS {eq_id}",
              f"#define CAT_EQ_DTHETA_SHIPPED {shipped['dtheta_dt'][0]}",
              f"#define CAT_EQ_DGAMMA_SHIPPED {shipped['dgamma_dt'][0]}",
              f"#define CAT_DTHETA_LIVE_MASK {shipped['dtheta_dt'][2]}u",
              f"#define CAT_DGAMMA_LIVE_MASK {shipped['dgamma_dt'][2]}u",
              "",
              "// shipped pair (eq_*.txt): the functions the fused MPC kernel inlines",
              "template <typename T, typename M>",
              "__device__ __forceinline__ T cat_dtheta_dt(const T (&x)[CAT_N_FEATURES]) { return "
              + shipped["dtheta_dt"][1] + "<T, M>(x); }",
              "template <typename T, typename M>",
              "__device__ __forceinline__ T cat_dgamma_dt(const T (&x)[CAT_N_FEATURES]) { return "
              + shipped["dgamma_dt"][1] + "<T, M>(x); }",
              "",
              "// N feature rows at once (the RK4 stages of one step): transcendental chains interleave",
              "template <typename T, typename M, int N>",
              "__device__ __forceinline__ void cat_dtheta_dt_v(const T (&x)[N][CAT_N_FEATURES], T (&out)[N]) { "
              + shipped["dtheta_dt"][1] + "_v<T, M, N>(x, out); }",
              "template <typename T, typename M, int N>",
              "__device__ __forceinline__ void cat_dgamma_dt_v(const T (&x)[N][CAT_N_FEATURES], T (&out)[N]) { "
              + shipped["dgamma_dt"][1] + "_v<T, M, N>(x, out); }",
              "",
              "// any hall-of-fame row by id (open-loop RK4 / predict kernels)",
              "template <typename T, typename M>",
              "__device__ __forceinline__ T cat_eval_equation(int eq_id, const T (&x)[CAT_N_FEATURES]) {",
              "  switch (eq_id) {"]
    for i, name, cx, mask, fn in table:
        lines.append(f"    case {i}: return {fn}<T, M>(x);")
    lines += ["    default: return T(0);", "  }", "}", "",
              "template <typename T, typename M, int N>",
              "__device__ __forceinline__ void cat_eval_equation_v(int eq_id, const T (&x)[N][CAT_N_FEATURES], T (&out)[N]) {",
              "  switch (eq_id) {"]
    for i, name, cx, mask, fn in table:
        lines.append(f"    case {i}: {fn}_v<T, M, N>(x, out); break;")
    lines += ["    default:", "#pragma unroll", "      for (int n = 0; n < N; ++n) out[n] = T(0);", "  }", "}", "",
              "struct CatEqInfo { int id; int is_gamma; int complexity; unsigned live_mask; };",
              "static const CatEqInfo CAT_EQ_TABLE[CAT_N_EQUATIONS] = {"]
    for i, name, cx, mask, fn in table:
        lines.append(f"  {{{i}, {1 if name == 'dgamma_dt' else 0}, {cx}, {mask}u}},")
    lines += ["};", ""]
    os.makedirs(os.path.dirname(out_file), exist_ok=True)
    generate_icdf()
    return _write_if_changed(out_file, "\n".join(lines))


def _write_if_changed(path, text):
    old = open(path).read() if os.path.exists(path) else None
    if old != text:
        with open(path, "w") as fh:
            fh.write(text)
    return path


def generate_icdf(table_json=ICDF_JSON, out_file=ICDF_FILE):
    """artifacts/icdf_table.json (tools/make_icdf_table.py) -> csrc/gen/icdf_table.cuh: the piecewise-cubic inverse
    normal CDF of the on-device control sampler, as one __device__ array of (c0, c1, c2, c3) rows."""
    with open(table_json) as fh:
        t = json.load(fh)
    coef = t["coef"]
    assert t["degree"] == 3 and t["sub_bits"] == 5 and len(coef) == (t["e_hi"] - t["e_lo"] + 1) << 5
    lines = ["// GENERATED by codegen.py from artifacts/icdf_table.json (tools/make_icdf_table.py) -- do not edit.",
             "// y(r') = -ndtri((r' + 0.5) / 2^32) on a floating-point-like grid: octave e = floor(log2 r'), 32 segments",
             "// per octave, cubic in the remaining bits.  max |error| " + repr(t["max_abs_err"]) + ".",
             "#pragma once",
             f"#define CAT_ICDF_E_LO {t['e_lo']}",
             f"#define CAT_ICDF_E_HI {t['e_hi']}",
             f"#define CAT_ICDF_SEGMENTS {len(coef)}",
             f"__device__ __align__(16) const double CAT_ICDF_TAB[{4 * len(coef)}] = {{"]
    for row in coef:
        lines.append("    " + ", ".join(repr(float(v)) for v in row) + ",")
    lines += ["};", ""]
    os.makedirs(os.path.dirname(out_file), exist_ok=True)
    return _write_if_changed(out_file, "\n".join(lines))


def equation_table(model_json=MODEL_JSON):
    """[(eq_id, name, complexity, sympy_format)] in the id order the generated header uses."""
    with open(model_json) as fh:
        model = json.load(fh)
    out = []
    for name in ("dtheta_dt", "dgamma_dt"):
        for row in model[name]["rows"]:
            out.append((len(out), name, row["complexity"], row["sympy_format"]))
    return out


def shipped_ids(model_json=MODEL_JSON):
    with open(model_json) as fh:
        model = json.load(fh)
    ids = {}
    for i, name, cx, _ in equation_table(model_json):
        if cx == model[name]["shipped_complexity"]:
            ids[name] = i
    return ids


if __name__ == "__main__":
    print("wrote", generate())
