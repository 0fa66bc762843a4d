"""NumPy/SciPy fp64 restatement of the reference's catenary + RK4 hot path.

TEST INFRASTRUCTURE -- the checker, never the product.  Citations are relative
to the reference checkout (eather0056/Catenary-Model-Estimation-and-MPC-Control-
for-ROV-Tethered-Systems).

Parity status
-------------
* PINNED (checked in tests/test_oracle_vs_reference_golden.py against fixtures
  produced by executing the reference's own importable functions, generator
  oracle/make_golden.py):
    rodrigues_rotation      main_fun.py:18-35
    transform_catenary      main_fun.py:38-111
    solve_catenary          main_fun.py:418-431  (scipy brentq, nan domain)
    extract_features        simple.py:15-41      (18-column layout)
    rk4_integration         simulate_rk4_theta_gamma.py:52-68
    integrate_theta_gamma   main_fun.py:735-764
    scaler / equations      saved_models/scaler.pkl, equations_*.csv
    compute_catenary_3D     models/catenary_3d.py:5-39
* PARITY UNPINNED: ``Catenary`` (pympc.models.catenary, git submodule pinned at
  f34f4c41bcb65f5a7291d3d519316c08d1693c7f, source absent, no URL).  The class
  here is the builder's definition: same call surface as the call sites
  (catenary.py:10,25; main_fun.py:64-69), physics of main_fun.py:423 and
  catenary_model.py:10-12.
* BUILDER-DEFINED (no reference counterpart): the closed-loop
  rollout, the cost and the sampling-MPC argmin.  This file IS their definition.
"""
from __future__ import annotations

import json
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_MODEL_JSON = os.path.join(
    _HERE, "..",
    "catenary-model-estimation-and-mpc-control-for-rov-tethered-systems_b200",
    "artifacts", "model.json")

# bracket of the reference brentq call, main_fun.py:425
C_LO = 1e-6
C_HI = 10.0
N_PTS = 100            # catenary_model.py:28, catenary_3d.py:52
NAN_PENALTY = 1e12     # SURVEY.md section 8a "nan C => +1e12"


def load_model():
    """Scaler constants and equation table extracted from saved_models/ (tools/extract_artifacts.py)."""
    with open(_MODEL_JSON) as fh:
        return json.load(fh)


# --------------------------------------------------------------------------
# geometry: main_fun.py:18-35, 38-111
# --------------------------------------------------------------------------
def rodrigues_rotation(vector, axis, angle_rad):
    """main_fun.py:18-35, batched over leading dims (last dim = 3)."""
    vector = np.asarray(vector, dtype=np.float64)
    axis = np.asarray(axis, dtype=np.float64)
    angle = np.asarray(angle_rad, dtype=np.float64)[..., None]
    axis = axis / np.linalg.norm(axis, axis=-1, keepdims=True)
    c = np.cos(angle)
    s = np.sin(angle)
    dot = np.sum(axis * vector, axis=-1, keepdims=True)
    return vector * c + np.cross(axis, vector) * s + axis * dot * (1 - c)


def theta_axis_of(conn):
    """Rotation axis for theta, main_fun.py:75-89 (fallbacks below 1e-9). conn (...,3)."""
    conn = np.asarray(conn, dtype=np.float64)
    xy = conn.copy()
    xy[..., 2] = 0.0
    n = np.linalg.norm(xy, axis=-1, keepdims=True)
    small = n < 1e-9
    xy = np.where(small, np.array([1.0, 0.0, 0.0]), xy / np.where(small, 1.0, n))
    ax = np.cross(xy, np.array([0.0, 0.0, 1.0]))
    n2 = np.linalg.norm(ax, axis=-1, keepdims=True)
    small2 = n2 < 1e-9
    return np.where(small2, np.array([0.0, 1.0, 0.0]), ax / np.where(small2, 1.0, n2))


def transform_catenary(point_A, point_B, catenary_fn, theta_rad, gamma_rad):
    """main_fun.py:38-111 for ONE configuration, vectorised over the curve points.

    catenary_fn(start, end) -> tuple whose [3] is (N,3) or None (then [start,end]).
    """
    point_A = np.asarray(point_A, dtype=np.float64)
    point_B = np.asarray(point_B, dtype=np.float64)

    def compute(start, end):
        out = catenary_fn(start, end)
        if out[3] is not None:
            return np.asarray(out[3], dtype=np.float64)
        return np.array([start, end])

    original = compute(point_A, point_B)
    conn = point_B - point_A
    k_theta = theta_axis_of(conn)
    rotated_B = point_A + rodrigues_rotation(conn, k_theta, theta_rad)
    theta_rotated = compute(point_A, rotated_B)
    theta_aligned = point_A + rodrigues_rotation(theta_rotated - point_A, k_theta, -theta_rad)
    k_gamma = conn / np.linalg.norm(conn)
    final = point_A + rodrigues_rotation(theta_aligned - point_A, k_gamma, gamma_rad)
    return original, theta_rotated, theta_aligned, final


def transform_catenary_batch(A, B, L, theta, gamma, n_pts=None):
    """transform_catenary (main_fun.py:38-111) with the stand-in Catenary, vectorised over n configurations:
    A, B (n,3); L, theta, gamma (n,) -> (original, theta_rotated, theta_aligned, final) each (n, n_pts, 3),
    C (n,) of the span A->B, valid (n,) bit 0: span A->B in the solve domain, bit 1: span A->B'.
    Same operations as the per-configuration function; a span outside the domain holds the straight chord
    sampled at n_pts points (its first and last samples are the [start, end] of main_fun.py:67-69)."""
    n_pts = N_PTS if n_pts is None else n_pts
    A = np.asarray(A, np.float64)
    B = np.asarray(B, np.float64)
    L = np.broadcast_to(np.asarray(L, np.float64), A.shape[:1])
    th = np.asarray(theta, np.float64)
    ga = np.asarray(gamma, np.float64)
    C0, v0, original = catenary_points(A, B, L, n_pts)
    conn = B - A
    k_theta = theta_axis_of(conn)
    rotated_B = A + rodrigues_rotation(conn, k_theta, th)
    _, v1, theta_rotated = catenary_points(A, rotated_B, L, n_pts)
    aligned = A[:, None] + rodrigues_rotation(theta_rotated - A[:, None], k_theta[:, None], -th[:, None])
    k_gamma = conn / np.linalg.norm(conn, axis=-1, keepdims=True)
    final = A[:, None] + rodrigues_rotation(aligned - A[:, None], k_gamma[:, None], ga[:, None])
    return (original, theta_rotated, aligned, final), C0, v0.astype(np.uint8) | (v1.astype(np.uint8) << 1)


# --------------------------------------------------------------------------
# catenary parameter: main_fun.py:418-431
# --------------------------------------------------------------------------
def _f_ref(C, l, dH, L):
    """main_fun.py:423, same operation order."""
    return C ** 2 * (L ** 2 - dH ** 2) - 4 * (np.sinh(0.5 * l * C)) ** 2


def solve_catenary_brentq(l, delta_H, L):
    """main_fun.py:418-431 restated: per-element scipy brentq on [1e-6, 10], nan on any failure.

    Slow (Python loop) -- the exact-semantics oracle for small cases.
    """
    from scipy.optimize import root_scalar
    l, delta_H, L = np.broadcast_arrays(np.asarray(l, np.float64), np.asarray(delta_H, np.float64),
                                        np.asarray(L, np.float64))
    out = np.empty(l.shape, dtype=np.float64)
    for idx in np.ndindex(l.shape):
        li, hi, Li = float(l[idx]), float(delta_H[idx]), float(L[idx])
        try:
            with np.errstate(all="ignore"):
                out[idx] = root_scalar(lambda C: _f_ref(C, li, hi, Li), bracket=[C_LO, C_HI],
                                       method="brentq").root
        except Exception:
            out[idx] = np.nan
    return out


def _sinh_minus_u(u):
    """sinh(u) - u without cancellation (series below 1)."""
    w = u * u
    ser = np.zeros_like(u)
    for k in range(10, 0, -1):           # u^3/3! (1 + w/(4*5) (1 + w/(6*7) ...))
        ser = 1.0 + ser * w / ((2 * k + 2) * (2 * k + 3))
    ser = ser * u * w / 6.0
    with np.errstate(over="ignore"):
        big = np.sinh(u) - u
    return np.where(np.abs(u) < 1.0, ser, big)


def _cosh_minus_1(u):
    with np.errstate(over="ignore"):
        return 2.0 * np.sinh(0.5 * u) ** 2


def solve_u(r_minus_1, iters=80):
    """Root u>0 of sinh(u)/u = 1 + e, e>0, vectorised safeguarded Newton run to convergence."""
    e = np.asarray(r_minus_1, dtype=np.float64)
    with np.errstate(all="ignore"):
        u_small = np.sqrt(6.0 * e / (1.0 + 0.3 * e))
        lr = np.log(2.0 * (1.0 + e))
        u_big = lr + np.log(lr + np.log(np.maximum(lr, 1.0)))
        u = np.where(e < 2.0, u_small, u_big)
        for _ in range(iters):
            g = _sinh_minus_u(u) - e * u
            gp = _cosh_minus_1(u) - e
            # g is convex for u>0: a Newton step from the right of the root is monotone;
            # from the left it jumps to the right.  Guard the flat region gp<=0.
            step = np.where(gp > 0, g / gp, -0.5 * u - 0.1)
            un = u - step
            un = np.where(un <= 0, 0.5 * u, un)
            done = np.all(np.abs(un - u) <= 4e-16 * np.abs(un))
            u = un
            if done:
                break
    return u


def solve_catenary(l, delta_H, L, c_lo=C_LO, c_hi=C_HI):
    """Vectorised equivalent of main_fun.py:418-431.

    nan domain reproduced with the reference's own sign test at the bracket ends
    (brentq raises when f(1e-6) and f(10) have the same sign -> nan, main_fun.py:427-428);
    the root itself comes from a converged Newton solve in u = C*l/2 (tighter than brentq's
    xtol=2e-12, so agreement with the reference is limited by brentq, ~1e-11 relative).
    """
    l, dH, L = np.broadcast_arrays(np.asarray(l, np.float64), np.asarray(delta_H, np.float64),
                                   np.asarray(L, np.float64))
    with np.errstate(all="ignore"):
        fa = _f_ref(c_lo, l, dH, L)
        fb = _f_ref(c_hi, l, dH, L)
        ok = ((fa > 0) & (fb < 0)) | ((fa < 0) & (fb > 0))
        # r - 1 = (sqrt(L^2-dH^2) - l)/l computed without cancellation
        s2 = (L - dH) * (L + dH)
        s = np.sqrt(np.where(s2 > 0, s2, np.nan))
        e = (s2 - l * l) / (l * (s + l))
        u = solve_u(np.where(ok & (e > 0), e, 1.0))
        C = 2.0 * u / l
    C = np.where(ok, C, np.nan)
    C = np.where(fa == 0, c_lo, C)
    C = np.where((fb == 0) & (fa != 0), c_hi, C)
    return C


def tension_from_C(C, l, w=1.521):
    """main_fun.py:302-305 -- T = w*l/(2*sinh(C*l/2)), nan -> w*l/2 (w: test_cluster.py:23)."""
    with np.errstate(all="ignore"):
        T = w * l / (2.0 * np.sinh(C * l / 2.0))
    return np.where(np.isnan(C), w * l / 2.0, T)


# --------------------------------------------------------------------------
# the missing pympc Catenary -- BUILDER-DEFINED stand-in (parity unpinned)
# --------------------------------------------------------------------------
def lowest_point_terms(l, dH, L, C, zA):
    """Vertex offset and lowest z (SURVEY.md section 8a 'per-step catenary terms').

    x_low = l/2 - atanh(dH/L)/C measured horizontally from A; z_low = zA + (1-cosh(C x_low))/C
    if 0 <= x_low <= l else min(zA, zB).  nan C -> min(zA, zB).
    """
    with np.errstate(all="ignore"):
        x_low = 0.5 * l - np.arctanh(dH / L) / C
        z_in = zA + (1.0 - np.cosh(C * x_low)) / C
    inside = (x_low >= 0) & (x_low <= l) & ~np.isnan(C)
    return x_low, np.where(inside, z_in, np.minimum(zA, zA + dH))


def catenary_points(A, B, L, n_pts=N_PTS, c_lo=C_LO, c_hi=C_HI):
    """Sampled hanging-cable curve between A and B (...,3): returns (C, valid, pts (...,n_pts,3)).

    Frame ENU: z up, cable sags toward -z (catenary.py:10,43).  Samples are uniform in the
    horizontal coordinate x_i = l*i/(n-1) (catenary_model.py:28) along the A->B azimuth;
    z_i = zA + (cosh(C (x_i - x_low)) - cosh(C x_low))/C  (catenary_model.py:10-12 shifted to the
    vertex of the unequal-height span).  valid == solve_catenary(...) is not nan.
    Invalid rows hold the straight chord (callers replace them by [A,B], main_fun.py:67-69).
    """
    A = np.asarray(A, np.float64)
    B = np.asarray(B, np.float64)
    rel = B - A
    l = np.hypot(rel[..., 0], rel[..., 1])
    dH = rel[..., 2]
    Lb = np.broadcast_to(np.asarray(L, np.float64), l.shape)
    C = solve_catenary(l, dH, Lb, c_lo, c_hi)
    valid = ~np.isnan(C)
    s = np.arange(n_pts, dtype=np.float64) / (n_pts - 1)
    with np.errstate(all="ignore"):
        x_low = 0.5 * l - np.arctanh(dH / Lb) / C
        x = l[..., None] * s
        z = A[..., 2:3] + (np.cosh(C[..., None] * (x - x_low[..., None])) -
                           np.cosh(C * x_low)[..., None]) / C[..., None]
    z_line = A[..., 2:3] + dH[..., None] * s
    z = np.where(valid[..., None], z, z_line)
    px = A[..., 0:1] + rel[..., 0:1] * s
    py = A[..., 1:2] + rel[..., 1:2] * s
    # pin the end points exactly
    pts = np.stack([px, py, z], axis=-1)
    return C, valid, pts


class Catenary:
    """Stand-in for pympc.models.catenary.Catenary (call sites catenary.py:10,25,51).

    ``Catenary(length=3., reference_frame='ENU')(a, b)`` returns a 4-tuple
    ``(C, H, D, points)``: catenary parameter C (1/m), sag H below the lower attachment
    point (m, >=0), horizontal distance D from ``a`` to the lowest point (m), and
    ``points`` = (N,3) array or None when the configuration is outside the solve domain.
    Only index 3 is ever read by the reference's callers.
    """

    def __init__(self, length=3.0, reference_frame="ENU", n_points=N_PTS):
        if reference_frame not in ("ENU", "NED"):
            raise ValueError("reference_frame must be 'ENU' or 'NED'")
        self.length = float(length)
        self.reference_frame = reference_frame
        self.n_points = int(n_points)

    def __call__(self, a, b):
        a = np.asarray(a, np.float64)
        b = np.asarray(b, np.float64)
        sgn = 1.0 if self.reference_frame == "ENU" else -1.0   # NED: z down, sag toward +z
        a2 = a * np.array([1.0, 1.0, sgn])
        b2 = b * np.array([1.0, 1.0, sgn])
        C, valid, pts = catenary_points(a2, b2, self.length, self.n_points)
        if not bool(valid):
            return None, None, None, None
        rel = b2 - a2
        l = float(np.hypot(rel[0], rel[1]))
        x_low, z_low = lowest_point_terms(l, rel[2], self.length, C, a2[2])
        H = float(min(a2[2], b2[2]) - z_low)
        D = float(min(max(x_low, 0.0), l))
        pts = pts * np.array([1.0, 1.0, sgn])
        return float(C), H, D, pts


def catenary_3d_parameter(p0, p1, rope_length):
    """The ``a`` compute_catenary_3D ends with (0 for a taut rope).  NOTE: the reference's fixed point
    a <- a L / (2 a sinh(d/2a)) is repelling for a slack rope -- started at d/2 it grows geometrically for the
    full 100 passes (a ~ 1e8 for d = 10, L = 12), so the sampled sag a cosh(d/2a) - a cosh(x/a) is formed with
    an absolute cancellation error of ~a * 2^-52.  Parity tolerances on this function are scaled with a."""
    d = np.asarray(p1, np.float64) - np.asarray(p0, np.float64)
    dist = np.sqrt(d[0] ** 2 + d[1] ** 2 + d[2] ** 2)
    if rope_length <= dist:
        return 0.0
    a = dist / 2
    for _ in range(100):
        a_new = a * rope_length / (2 * a * np.sinh(dist / (2 * a)))
        if abs(a_new - a) < 1e-6:
            return float(a_new)
        a = a_new
    return float(a)


def compute_catenary_3D(p0, p1, rope_length, num_points):
    """models/catenary_3d.py:5-39 restated (fixed-point in a, symmetric sag profile)."""
    p0 = np.asarray(p0, np.float64)
    p1 = np.asarray(p1, np.float64)
    d = p1 - p0
    dist = np.sqrt(d[0] ** 2 + d[1] ** 2 + d[2] ** 2)
    if rope_length <= dist:
        return np.linspace(p0, p1, num_points)
    half = dist / 2
    a = half
    for _ in range(100):
        lhs = 2 * a * np.sinh(dist / (2 * a))
        a_new = a * rope_length / lhs
        if abs(a_new - a) < 1e-6:
            a = a_new
            break
        a = a_new
    off = a * np.cosh(half / a)
    t = np.arange(num_points) / (num_points - 1)
    pts = p0[None, :] + d[None, :] * t[:, None]
    sag = off - a * np.cosh((t * dist - half) / a)
    pts[:, 2] -= sag
    return pts


# --------------------------------------------------------------------------
# features / scaler / equations: simple.py:15-41, saved_models/*
# --------------------------------------------------------------------------
def features18(P0, P1, V1, A1, theta, gamma, theta_prev, gamma_prev):
    """Row-wise 18-column feature layout of simple.py:26-41 from already-built columns.

    P0,P1 in m; V1 in mm/s (simple.py:16-18 leaves velocities unscaled); A1 in mm/s^2.
    """
    P0, P1, V1, A1 = (np.asarray(a, np.float64) for a in (P0, P1, V1, A1))
    rel = P1 - P0
    nrm = np.linalg.norm(rel, axis=-1, keepdims=True)
    unit = rel / (nrm + 1e-8)
    tension = np.clip(nrm, 1e-5, 10)
    dot = np.sum(V1 * unit, axis=-1, keepdims=True)
    nv = np.linalg.norm(V1, axis=-1, keepdims=True) + 1e-8
    ang = np.clip(dot / nv, -1, 1)
    col = lambda a: np.asarray(a, np.float64)[..., None]
    return np.concatenate([P1, V1, A1, unit, tension, ang, col(theta), col(gamma),
                           col(theta_prev), col(gamma_prev)], axis=-1)


def extract_features(df):
    """simple.py:15-41 restated for a DataFrame (or dict of columns) in data units (mm, mm/s)."""
    g = lambda k: np.asarray(df[k], np.float64)
    P0 = np.stack([g("rod_end X"), g("rod_end Y"), g("rod_end Z")], axis=1) / 1000
    P1 = np.stack([g("robot_cable_attach_point X"), g("robot_cable_attach_point Y"),
                   g("robot_cable_attach_point Z")], axis=1) / 1000
    V1 = np.stack([g("rob_cor_speed X"), g("rob_cor_speed Y"), g("rob_cor_speed Z")], axis=1)
    t = g("Time")
    A1 = np.stack([np.gradient(V1[:, j], t) for j in range(3)], axis=1)
    th = g("Theta")
    ga = g("Gamma")
    thp = np.roll(th, 1)
    gap = np.roll(ga, 1)
    thp[0] = th[0]
    gap[0] = ga[0]
    return features18(P0, P1, V1, A1, th, ga, thp, gap)


def scale_features(X, model=None):
    """StandardScaler.transform with saved_models/scaler.pkl constants (test_temp.py:49)."""
    m = model or load_model()
    return (np.asarray(X, np.float64) - np.array(m["scaler_mean"])) / np.array(m["scaler_scale"])


_EQ_CACHE = {}


def equation(name, complexity=None, model=None):
    """NumPy callable f(X (...,18)) for a hall-of-fame row, built like PySR's predict does:
    sympy-lambdify of the CSV ``sympy_format`` string (equations_d*_dt.csv)."""
    m = model or load_model()
    entry = m[name]
    cx = entry["shipped_complexity"] if complexity is None else complexity
    key = (name, cx)
    if key not in _EQ_CACHE:
        import sympy
        row = next(r for r in entry["rows"] if r["complexity"] == cx)
        syms = sympy.symbols("x0:18")
        expr = sympy.sympify(row["sympy_format"], locals={f"x{i}": s for i, s in enumerate(syms)})
        fn = sympy.lambdify(syms, expr, "numpy")

        def f(X, _fn=fn):
            X = np.asarray(X, np.float64)
            out = _fn(*[X[..., i] for i in range(18)])
            return np.broadcast_to(np.asarray(out, np.float64), X.shape[:-1]).copy()
        _EQ_CACHE[key] = f
    return _EQ_CACHE[key]


class EquationModel:
    """Minimal object with the ``predict`` surface rk4_integration / integrate_theta_gamma need."""

    def __init__(self, name, complexity=None):
        self.name = name
        self.f = equation(name, complexity)

    def predict(self, X):
        return self.f(np.asarray(X, np.float64))


# --------------------------------------------------------------------------
# integrators: simulate_rk4_theta_gamma.py:52-68, main_fun.py:735-764
# --------------------------------------------------------------------------
def rk4_integration(model, x_input, time, y0):
    """simulate_rk4_theta_gamma.py:52-68 restated (open loop: y never feeds f; k2 == k3)."""
    x_input = np.asarray(x_input, np.float64)
    y = [y0]
    for i in range(1, len(time)):
        dt = time[i] - time[i - 1]
        x0 = x_input[i - 1]
        x1 = x_input[i]
        f = lambda x: model.predict(x.reshape(1, -1))[0]
        k1 = f(x0)
        k2 = f((x0 + x1) / 2)
        k3 = f((x0 + x1) / 2)
        k4 = f(x1)
        y.append(y[-1] + (dt / 6) * (k1 + 2 * k2 + 2 * k3 + k4))
    return np.array(y)


def rk4_integration_vec(model, x_input, time, y0):
    """Same arithmetic as rk4_integration, stage evaluations vectorised over time, the
    accumulation kept sequential (identical association)."""
    x = np.asarray(x_input, np.float64)
    time = np.asarray(time, np.float64)
    if len(time) == 0:
        return np.zeros(0)
    k1 = model.predict(x[:-1])
    k2 = model.predict((x[:-1] + x[1:]) / 2)
    k4 = model.predict(x[1:])
    inc = (np.diff(time) / 6) * (k1 + 2 * k2 + 2 * k2 + k4)
    y = np.empty(len(time))
    y[0] = y0
    for i in range(1, len(time)):
        y[i] = y[i - 1] + inc[i - 1]
    return y


def integrate_theta_gamma(model_theta, model_gamma, X, time_array, theta_0, gamma_0):
    """main_fun.py:735-764 restated (explicit Euler, open loop)."""
    n = len(time_array)
    th = np.zeros(n)
    ga = np.zeros(n)
    th[0] = theta_0
    ga[0] = gamma_0
    for i in range(1, n):
        dt = time_array[i] - time_array[i - 1]
        th[i] = th[i - 1] + model_theta.predict(X[i - 1:i])[0] * dt
        ga[i] = ga[i - 1] + model_gamma.predict(X[i - 1:i])[0] * dt
    return th, ga


# --------------------------------------------------------------------------
# BUILDER-DEFINED: closed-loop rollout + cost + argmin
# --------------------------------------------------------------------------
def default_mpc_params(**over):
    """Config-3 constants: pose of fully_augmented_catenary.py:11-15."""
    p = dict(
        dt=0.01,                                   # see DESIGN.md "closed-loop dt" (main_fun.py:127 uses 0.05 open loop)
        L=3.0,                                     # catenary.py:10
        P0=(0.0, 0.0, 0.0),                        # catenary.py:13
        P1=(1.0, 1.0, -1.0),                       # fully_augmented_catenary.py:12
        V1=(0.0, 0.0, 0.0),
        A1=(0.0, 0.0, 0.0),
        theta=float(np.radians(-10.0)),            # fully_augmented_catenary.py:14
        gamma=float(np.radians(20.0)),             # fully_augmented_catenary.py:15
        theta_prev=float(np.radians(-10.0)),
        gamma_prev=float(np.radians(20.0)),
        w_trk=1.0, w_du=1e-4, w_len=1e3, w_sag=1e3, w_ang=0.1,
        rho=0.95, z_floor=-2.0, margin=0.1,
        c_lo=C_LO, c_hi=C_HI, nan_penalty=NAN_PENALTY,
        theta_cx=None, gamma_cx=None,              # hall-of-fame rows; None = shipped
        theta_expr=None, gamma_expr=None,          # run-time equation strings over x0..x17 (override the rows)
    )
    p.update(over)
    return p


def circle_reference(H, dt, center=(1.0, 1.0, -1.0), radius=0.1, period=10.0):
    """Target P_ref[h] for h=1..H: 'case 12' circular path of Rov_traj_gen.py:113-117 (r=0.1 m,
    period 10 s) passing through the initial pose at t=0."""
    t = (np.arange(H) + 1) * dt
    c = np.asarray(center, np.float64)
    w = 2 * np.pi / period
    return np.stack([c[0] - radius + radius * np.cos(w * t),
                     c[1] + radius * np.sin(w * t),
                     np.full(H, c[2])], axis=1)


def mpc_rollout_costs(params, controls, p_ref, return_traj=False):
    """Cost J[k] of every candidate control sequence.

    controls (K,H,3): commanded V1 in mm/s; p_ref (H,3) in m.  Per step h (all K at once):
      Vn = u_h;  Pn = P + Vn*dt/1000;  An = (Vn - V)/dt
      X0 = scale(feat(P ,V ,A ,th    ,ga    ,th_prev,ga_prev))
      k1 = f(X0);  th_hat = th + dt*k1_th;  ga_hat = ga + dt*k1_ga
      X1 = scale(feat(Pn,Vn,An,th_hat,ga_hat,th     ,ga     ))
      k2 = k3 = f((X0+X1)/2);  k4 = f(X1)            (simulate_rk4_theta_gamma.py:61-64)
      y  += dt/6*(k1 + 2k2 + 2k3 + k4)               (simulate_rk4_theta_gamma.py:66)
      l,dH,d from Pn-P0;  C = solve_catenary(l,dH,L)  (main_fun.py:418-431)
      z_low from lowest_point_terms
      c_h = ((((w_trk|Pn-Pref_h|^2 + w_du|Vn-V|^2) + w_len max(0,d-rho L)^2)
               + w_sag max(0,z_floor+margin-z_low)^2) + w_ang(th^2+ga^2)) [+ nan_penalty if C nan]
      J += c_h
    """
    p = params
    m = load_model()
    mean = np.array(m["scaler_mean"])
    inv = 1.0 / np.array(m["scaler_scale"])
    if p.get("theta_expr") is not None:             # run-time equation strings
        f_th = LambdifiedModel(p["theta_expr"], 18).predict
        f_ga = LambdifiedModel(p["gamma_expr"], 18).predict
    else:
        f_th = equation("dtheta_dt", p.get("theta_cx"), m)
        f_ga = equation("dgamma_dt", p.get("gamma_cx"), m)
    U = np.asarray(controls, np.float64)
    K, H, _ = U.shape
    p_ref = np.asarray(p_ref, np.float64)
    dt = float(p["dt"])
    L = float(p["L"])
    P0 = np.asarray(p["P0"], np.float64)
    P = np.broadcast_to(np.asarray(p["P1"], np.float64), (K, 3)).copy()
    V = np.broadcast_to(np.asarray(p["V1"], np.float64), (K, 3)).copy()
    A = np.broadcast_to(np.asarray(p["A1"], np.float64), (K, 3)).copy()
    th = np.full(K, float(p["theta"]))
    ga = np.full(K, float(p["gamma"]))
    thp = np.full(K, float(p["theta_prev"]))
    gap = np.full(K, float(p["gamma_prev"]))
    J = np.zeros(K)
    traj = {k: [] for k in ("P", "theta", "gamma", "C", "z_low", "step_cost")}
    scale = lambda X: (X - mean) * inv
    for h in range(H):
        Vn = U[:, h, :]
        Pn = P + Vn * (dt / 1000.0)
        An = (Vn - V) / dt
        X0 = scale(features18(P0, P, V, A, th, ga, thp, gap))
        k1t = f_th(X0)
        k1g = f_ga(X0)
        X1 = scale(features18(P0, Pn, Vn, An, th + dt * k1t, ga + dt * k1g, th, ga))
        Xm = (X0 + X1) / 2
        k2t = f_th(Xm)
        k2g = f_ga(Xm)
        k4t = f_th(X1)
        k4g = f_ga(X1)
        thn = th + (dt / 6) * (k1t + 2 * k2t + 2 * k2t + k4t)
        gan = ga + (dt / 6) * (k1g + 2 * k2g + 2 * k2g + k4g)
        rel = Pn - P0
        l = np.hypot(rel[:, 0], rel[:, 1])
        dH = rel[:, 2]
        d = np.sqrt(rel[:, 0] ** 2 + rel[:, 1] ** 2 + rel[:, 2] ** 2)
        C = solve_catenary(l, dH, L, p["c_lo"], p["c_hi"])
        _, z_low = lowest_point_terms(l, dH, L, C, P0[2])
        e = Pn - p_ref[h]
        dv = Vn - V
        c = p["w_trk"] * (e[:, 0] ** 2 + e[:, 1] ** 2 + e[:, 2] ** 2)
        c = c + p["w_du"] * (dv[:, 0] ** 2 + dv[:, 1] ** 2 + dv[:, 2] ** 2)
        c = c + p["w_len"] * np.maximum(0.0, d - p["rho"] * L) ** 2
        c = c + p["w_sag"] * np.maximum(0.0, p["z_floor"] + p["margin"] - z_low) ** 2
        c = c + p["w_ang"] * (thn ** 2 + gan ** 2)
        c = np.where(np.isnan(C), c + p["nan_penalty"], c)
        J = J + c
        thp, gap, th, ga = th, ga, thn, gan
        P, V, A = Pn, Vn, An
        if return_traj:
            for k, v in (("P", P), ("theta", th), ("gamma", ga), ("C", C), ("z_low", z_low),
                         ("step_cost", c)):
                traj[k].append(np.array(v))
    if return_traj:
        return J, {k: np.stack(v, axis=1) for k, v in traj.items()}
    return J


def argmin_first(costs):
    """np.argmin semantics (fully_augmented_catenary.py:21, the curve's lowest sample): first minimum; first nan
    wins.  The MPC winner uses argmin_mpc."""
    costs = np.asarray(costs)
    i = int(np.argmin(costs))
    return i, float(costs[i])


def argmin_mpc(costs):
    """Winner of a sampling-MPC step (builder-defined): lowest cost, ties -> lowest index, and a NaN cost ranks
    with +inf -- a candidate whose rollout failed must never be applied to the plant.  With every cost NaN the
    lowest index wins (candidate 0 = the noise-free nominal sequence in the controller).  Returns (idx, cost)."""
    costs = np.asarray(costs, np.float64)
    i = int(np.argmin(np.where(np.isnan(costs), np.inf, costs)))
    return i, float(costs[i])


def mpc_sample_step(params, controls, p_ref):
    J = mpc_rollout_costs(params, controls, p_ref)
    i, c = argmin_mpc(J)
    return i, c, J


# --------------------------------------------------------------------------
# counter-based RNG for on-device control sampling (configs 4/5): Philox4x32-10
# --------------------------------------------------------------------------
_PH_M0 = np.uint64(0xD2511F53)
_PH_M1 = np.uint64(0xCD9E8D57)
_PH_W0 = np.uint32(0x9E3779B9)
_PH_W1 = np.uint32(0xBB67AE85)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Philox4x32 with 10 rounds (Salmon et al. 2011), vectorised over uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(c, np.uint32).copy() for c in np.broadcast_arrays(c0, c1, c2, c3))
    k0 = np.uint32(k0)
    k1 = np.uint32(k1)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = _PH_M0 * c0.astype(np.uint64)
            p1 = _PH_M1 * c2.astype(np.uint64)
            hi0 = (p0 >> np.uint64(32)).astype(np.uint32)
            lo0 = p0.astype(np.uint32)
            hi1 = (p1 >> np.uint64(32)).astype(np.uint32)
            lo1 = p1.astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0 = np.uint32((int(k0) + int(_PH_W0)) & 0xFFFFFFFF)
            k1 = np.uint32((int(k1) + int(_PH_W1)) & 0xFFFFFFFF)
    return c0, c1, c2, c3


_ICDF = None


def icdf_table():
    """artifacts/icdf_table.json (tools/make_icdf_table.py): the piecewise-cubic inverse normal CDF that DEFINES the
    sampler -- the CUDA kernels compile the same numbers (csrc/gen/icdf_table.cuh)."""
    global _ICDF
    if _ICDF is None:
        with open(os.path.join(os.path.dirname(_MODEL_JSON), "icdf_table.json")) as fh:
            t = json.load(fh)
        _ICDF = (int(t["e_lo"]), int(t["sub_bits"]), np.array(t["coef"], np.float64))
    return _ICDF


def icdf_normal(r):
    """Standard normal from uint32 words: z = sign * y(r'), r' = r if r < 2^31 else ~r (clamped to >= 32),
    sign = -1 for r < 2^31; y = cubic of the segment (octave e = floor(log2 r'), next five bits) in
    t = (bits below) / 2^(e-5)."""
    e_lo, m, coef = icdf_table()
    r = np.asarray(r, np.uint32)
    hi = r >= np.uint32(0x80000000)
    rp = np.where(hi, ~r, r).astype(np.int64)
    rp = np.maximum(rp, 32)
    e = np.floor(np.log2(rp.astype(np.float64))).astype(np.int64)      # exact for integers < 2^53
    w = np.left_shift(np.int64(1), e - m)
    sub = (rp >> (e - m)) & ((1 << m) - 1)
    t = (rp & (w - 1)).astype(np.float64) / w.astype(np.float64)
    c = coef[((e - e_lo) << m) + sub]
    y = ((c[..., 3] * t + c[..., 2]) * t + c[..., 1]) * t + c[..., 0]
    return np.where(hi, y, -y)


def sample_controls(seed, k_index, H, sigma, clip=3.0, problem=0):
    """controls[k,h,:] = sigma * clip(N(0,1), +-clip) keyed by (seed, problem, global k, h).

    counter = (k_lo, k_hi, h, problem), key = (seed_lo, seed_hi); Philox words r0, r1, r2 -> axes x, y, z through
    the tabulated inverse normal CDF (icdf_normal).
    """
    k_index = np.asarray(k_index, np.uint64)
    klo = (k_index & np.uint64(0xFFFFFFFF)).astype(np.uint32)[:, None]
    khi = (k_index >> np.uint64(32)).astype(np.uint32)[:, None]
    hh = np.arange(H, dtype=np.uint32)[None, :]
    r0, r1, r2, _ = philox4x32_10(klo, khi, hh, np.uint32(problem),
                                  seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    z = np.stack([icdf_normal(r0), icdf_normal(r1), icdf_normal(r2)], axis=-1)
    z = np.clip(z, -clip, clip)
    return z * np.asarray(sigma, np.float64)


# --------------------------------------------------------------------------
# BUILDER-DEFINED: receding-horizon controller around the sampling step
# --------------------------------------------------------------------------
CTRL_LOG_W = 16


def ctrl_candidates(u_nom, seed, t, K, sigma, u_max, clip=3.0, problem=0):
    """(K,H,3) candidates of control step t: clamp(u_nom[h] + sigma*clip(N(0,1)), +-u_max), Philox keyed by
    (seed + t, problem, k, h); candidate 0 carries no noise."""
    H = u_nom.shape[0]
    Z = sample_controls((int(seed) + int(t)) & 0xFFFFFFFFFFFFFFFF, np.arange(K), H, sigma, clip, problem)
    Z[0] = 0.0
    um = np.asarray(u_max, np.float64)
    return np.minimum(np.maximum(u_nom[None] + Z, -um), um)


def mpc_receding_horizon(params, p_ref_traj, n_steps, K, H, sigma, u_max, seed=0, mode="argmin", lam=1.0,
                         clip=3.0, u_nom0=None, problem=0):
    """Closed loop of n_steps control steps (plant = the model).  Returns dict(log (n_steps,16), u_nom (H,3),
    state (params dict after the last step)).  Per control step t:
      U      = ctrl_candidates(u_nom, seed, t, ...)
      window = p_ref_traj[min(t + h, ref_len - 1)], h = 0..H-1
      J      = mpc_rollout_costs(state, U, window);  i* = argmin_first(J)
      u*     = U[i*]                                   (mode 'argmin')
             = sum_k w_k U_k / sum_k w_k, w_k = exp(-(J_k - J_i*)/lam), nan J -> w = 0   (mode 'mppi';
               falls back to U[i*] when J_i* is nan)
      plant: one rollout step with u*[0] -> new (P1, V1, A1, theta, gamma, prevs)
      u_nom  = (u*[1], ..., u*[H-1], u*[H-1])
    """
    p = dict(params)
    traj = np.asarray(p_ref_traj, np.float64)
    ref_len = traj.shape[0]
    u_nom = np.zeros((H, 3)) if u_nom0 is None else np.array(u_nom0, np.float64)
    log = np.zeros((n_steps, CTRL_LOG_W))
    for t in range(n_steps):
        U = ctrl_candidates(u_nom, seed, t, K, sigma, u_max, clip, problem)
        win = traj[np.minimum(t + np.arange(H), ref_len - 1)]
        J = mpc_rollout_costs(p, U, win)
        i, Jb = argmin_mpc(J)
        if mode == "mppi" and not np.isnan(Jb):
            w = np.exp(-(J - Jb) / lam)
            w = np.where(np.isnan(J), 0.0, w)
            u_star = (w[:, None, None] * U).sum(axis=0) / w.sum()
        else:
            u_star = U[i].copy()
        c, tr = mpc_rollout_costs(p, u_star[None, :1], win[:1], return_traj=True)
        V_old = np.asarray(p["V1"], np.float64)
        P_new = tr["P"][0, 0]
        p["A1"] = tuple((u_star[0] - V_old) / float(p["dt"]))
        p["V1"] = tuple(u_star[0])
        p["P1"] = tuple(P_new)
        p["theta_prev"], p["gamma_prev"] = p["theta"], p["gamma"]
        p["theta"], p["gamma"] = float(tr["theta"][0, 0]), float(tr["gamma"][0, 0])
        rel = P_new - np.asarray(p["P0"], np.float64)
        log[t] = [*P_new, *u_star[0], p["theta"], p["gamma"], tr["C"][0, 0], tr["z_low"][0, 0], Jb, i,
                  c[0], t, np.sqrt(rel[0] ** 2 + rel[1] ** 2 + rel[2] ** 2), 0.0]
        u_nom = np.concatenate([u_star[1:], u_star[-1:]], axis=0)
    return dict(log=log, u_nom=u_nom, state=p)


# --------------------------------------------------------------------------
# run-time equations: checker for expr.py / cat_kernels_expr.cuh
# --------------------------------------------------------------------------
class LambdifiedModel:
    """predict() of an arbitrary equation string the way PySRRegressor does it: sympy.sympify + lambdify to
    NumPy (simple.py:92,105 wrote the sympy_format strings)."""

    def __init__(self, expr, n_features, feature_names=None, scaler_mean=None, scaler_scale=None):
        import sympy
        names = list(feature_names) if feature_names else [f"x{j}" for j in range(n_features)]
        syms = [sympy.Symbol(n) for n in names]
        loc = dict(zip(names, syms))
        loc.update(square=lambda v: v ** 2, cube=lambda v: v ** 3, neg=lambda v: -v)
        self.fn = sympy.lambdify(syms, sympy.sympify(str(expr).replace("^", "**"), locals=loc), modules="numpy")
        self.n = n_features
        self.mean = None if scaler_mean is None else np.asarray(scaler_mean, np.float64)
        self.scale = None if scaler_scale is None else np.asarray(scaler_scale, np.float64)

    def predict(self, X):
        X = np.asarray(X, np.float64)
        if self.mean is not None:
            X = (X - self.mean) / self.scale
        with np.errstate(all="ignore"):
            y = self.fn(*[X[:, j] for j in range(self.n)])
        return np.broadcast_to(np.asarray(y, np.float64), (X.shape[0],)).copy()


def integrate_second_order(pred, time, y0):
    """test_cluster.py:110-129 restated: the model predicts the SECOND derivative; rate from 0, angle from y0,
    both explicit Euler on the previous sample."""
    n = len(time)
    rate = np.zeros(n)
    y = np.zeros(n)
    y[0] = y0
    for i in range(1, n):
        dt = time[i] - time[i - 1]
        rate[i] = rate[i - 1] + pred[i - 1] * dt
    for i in range(1, n):
        dt = time[i] - time[i - 1]
        y[i] = y[i - 1] + rate[i - 1] * dt
    return y, rate


# --------------------------------------------------------------------------
# other feature layouts + resampler: checker for features.py
# --------------------------------------------------------------------------
def _experiment_arrays(df):
    g = lambda k: np.asarray(df[k], np.float64)
    P0 = np.stack([g("rod_end X"), g("rod_end Y"), g("rod_end Z")], 1) / 1000
    P1 = np.stack([g("robot_cable_attach_point X"), g("robot_cable_attach_point Y"),
                   g("robot_cable_attach_point Z")], 1) / 1000
    V1 = np.stack([g("rob_cor_speed X"), g("rob_cor_speed Y"), g("rob_cor_speed Z")], 1)
    t = g("Time")
    A1 = np.stack([np.gradient(V1[:, j], t) for j in range(3)], 1)
    return P0, P1, V1, A1, t, g("Theta"), g("Gamma")


def feature_terms(df, L=3.0, cable_wet_weight=1.521):
    """Every per-row term of main_fun.py:167-415 restated (same operations, same association)."""
    P0, P1, V1, A1, t, th, ga = _experiment_arrays(df)
    rel = P1 - P0
    nrm = np.linalg.norm(rel, axis=1, keepdims=True)
    unit = rel / (nrm + 1e-8)                                   # main_fun.py:283
    v_surge = np.sum(V1 * unit, axis=1, keepdims=True)          # :286
    v_sway = np.linalg.norm(np.cross(V1, unit), axis=1, keepdims=True)   # :287
    l = np.linalg.norm(rel[:, :2], axis=1, keepdims=True)       # :290
    dH = rel[:, 2].reshape(-1, 1)                               # :291
    wl = cable_wet_weight / L                                   # :300
    C = solve_catenary(l.ravel(), dH.ravel(), L).reshape(-1, 1)  # :301
    with np.errstate(all="ignore"):
        T = (wl * l) / (2 * np.sinh(C * l / 2))                 # :302
    T = np.where(np.isnan(T), wl * l / 2, T)                    # :303
    v1n = np.linalg.norm(V1, axis=1, keepdims=True)
    ap = np.clip(v_surge / (v1n + 1e-8), -1, 1)                 # :313
    a_dot = np.sum(A1 * unit, axis=1, keepdims=True)
    a_perp = A1 - a_dot * unit                                  # :317-318
    a1n = np.linalg.norm(A1, axis=1, keepdims=True)
    return dict(P1=P1, V1=V1, A1=A1, unit_rel=unit, dist_clip=np.clip(nrm, 1e-5, 10), angle_proj=ap,
                v_surge=v_surge, v_sway=v_sway, l=l, delta_H=dH, C=C, T=T, angle_proj_tanh=np.tanh(ap),
                a1_dot_unitrel=a_dot, a1_perp_norm=np.linalg.norm(a_perp, axis=1, keepdims=True), v1_norm=v1n,
                a1_norm=a1n, acc_vel_ratio=a1n / (v1n + 1e-8), theta=th.reshape(-1, 1), gamma=ga.reshape(-1, 1))


def build_theta_features(df, L, cable_wet_weight):
    """main_fun.py:267-340."""
    f = feature_terms(df, L, cable_wet_weight)
    return np.hstack([f["theta"], f["v_surge"], f["v_sway"], f["l"], f["delta_H"], f["T"], f["angle_proj_tanh"],
                      f["A1"][:, 1:2], f["a1_perp_norm"]])


def build_gamma_features(df, L, cable_wet_weight):
    """main_fun.py:344-415."""
    f = feature_terms(df, L, cable_wet_weight)
    return np.hstack([f["gamma"], f["v_sway"], f["v_surge"], f["l"], f["delta_H"], f["T"], f["angle_proj_tanh"],
                      f["A1"][:, 1:2], f["a1_perp_norm"]])


def build_clean_features(df):
    """main_fun.py:197-265."""
    f = feature_terms(df)
    return np.hstack([f["v_surge"], f["a1_dot_unitrel"], f["v_sway"], f["a1_perp_norm"], f["v1_norm"], f["a1_norm"],
                      f["acc_vel_ratio"], f["v1_norm"] ** 2, f["a1_norm"] ** 2, f["dist_clip"], f["angle_proj_tanh"],
                      f["V1"][:, 1:2], f["A1"][:, 1:2]])


def extract_features16(df):
    """main_fun.py:167-193."""
    f = feature_terms(df)
    return np.hstack([f["P1"], f["V1"], f["A1"], f["unit_rel"], f["dist_clip"], f["angle_proj"], f["theta"],
                      f["gamma"]])


def resample_linear(time_in, values, time_out):
    """scipy interp1d(kind='linear', fill_value='extrapolate') restated (main_fun.py:139-140; scipy 1.18
    interp1d._call_linear): searchsorted, clip to [1, n-1], ((x_new-x_lo)/(x_hi-x_lo)) y_hi + ((x_hi-x_new)/(x_hi-x_lo)) y_lo."""
    x = np.asarray(time_in, np.float64)
    y = np.asarray(values, np.float64)
    xn = np.asarray(time_out, np.float64)
    idx = np.clip(np.searchsorted(x, xn), 1, len(x) - 1)
    lo, hi = idx - 1, idx
    return ((xn - x[lo]) / (x[hi] - x[lo]))[:, None] * y[hi] + ((x[hi] - xn) / (x[hi] - x[lo]))[:, None] * y[lo]


# --------------------------------------------------------------------------
# smoothed / second-order feature layouts (second half): main_fun.py:504-655, 768-776, 811-871.
# The SciPy filters the reference calls are restated in NumPy (the GPU path follows these restatements); the golden
# vectors in tests/golden/feature_layouts2.npz come from the reference's own functions with the real SciPy.
# --------------------------------------------------------------------------
def gaussian_taps(sigma, truncate=4.0):
    """scipy.ndimage._gaussian_kernel1d(sigma, 0, radius) with radius = int(truncate * sigma + 0.5)."""
    radius = int(truncate * float(sigma) + 0.5)
    x = np.arange(-radius, radius + 1)
    phi = np.exp(-0.5 / (float(sigma) * float(sigma)) * x ** 2)
    return phi / phi.sum()


def fir_reflect(x, taps):
    """Correlation of the LAST axis with symmetric ``taps``, scipy.ndimage mode='reflect' (d c b a | a b c d | d c b a):
    what gaussian_filter1d(x, sigma) does (main_fun.py:512-519, 772-773)."""
    x = np.asarray(x, np.float64)
    n = x.shape[-1]
    r = (len(taps) - 1) // 2
    idx = np.arange(-r, n + r)
    idx = np.mod(idx, 2 * n)
    idx = np.where(idx >= n, 2 * n - 1 - idx, idx)
    ext = x[..., idx]
    out = np.zeros_like(x)
    for j, w in enumerate(taps):
        out = out + w * ext[..., j:j + n]
    return out


def gaussian_filter1d(x, sigma):
    return fir_reflect(x, gaussian_taps(sigma))


def savgol_tables(window=11, order=3):
    """(taps (window,), edge_left (h, window), edge_right (h, window)) of scipy.signal.savgol_filter(x, window, order)
    with its default mode='interp': interior = least-squares polynomial value at the window centre; the first / last
    h = window // 2 samples = the polynomial fitted to the first / last `window` samples, evaluated in place."""
    h = window // 2
    xs = np.arange(-h, h + 1, dtype=np.float64)
    taps = np.linalg.pinv(np.vander(xs, order + 1, increasing=True))[0]
    pos = np.arange(window, dtype=np.float64)
    P = np.linalg.pinv(np.vander(pos, order + 1, increasing=True))            # (order+1, window)
    edge_l = np.vander(pos[:h], order + 1, increasing=True) @ P
    edge_r = np.vander(pos[window - h:], order + 1, increasing=True) @ P
    return taps, edge_l, edge_r


def savgol_filter(x, window=11, order=3):
    x = np.asarray(x, np.float64)
    n = x.shape[-1]
    if n < window:
        raise ValueError("savgol_filter: mode='interp' needs window_length <= size of x")
    taps, el, er = savgol_tables(window, order)
    h = window // 2
    out = np.empty_like(x)
    for i in range(h, n - h):
        out[..., i] = x[..., i - h:i + h + 1] @ taps
    out[..., :h] = x[..., :window] @ el.T
    out[..., n - h:] = x[..., n - window:] @ er.T
    return out


def _valid_terms(df):
    P0, P1, V1, _, t, th, ga = _experiment_arrays(df)
    rel = P1 - P0
    l = np.linalg.norm(rel[:, :2], axis=1, keepdims=True)
    unit = rel / (np.linalg.norm(rel, axis=1, keepdims=True) + 1e-8)
    v_surge = np.sum(V1 * unit, axis=1, keepdims=True)
    v_sway = np.linalg.norm(np.cross(V1, unit), axis=1, keepdims=True)
    # gaussian_filter1d of a (T,1) column filters along the LAST axis -- length 1 -- so the reference's `theta` and
    # `gamma` feature columns are the raw signals times the weight sum (main_fun.py:510-511, 578-579) ...
    theta = gaussian_filter1d(th.reshape(-1, 1), 2)
    gamma = gaussian_filter1d(ga.reshape(-1, 1), 2)
    # ... while the derivatives are taken of the properly smoothed 1-D signals (main_fun.py:517-521)
    dtheta = np.gradient(gaussian_filter1d(th, 2), t).reshape(-1, 1)
    dgamma = np.gradient(gaussian_filter1d(ga, 2), t).reshape(-1, 1)
    return l, v_surge, v_sway, theta, gamma, dtheta, dgamma


def build_theta_features_valid(df, L=3.0, cable_wet_weight=1.521):
    """main_fun.py:504-571 (13 columns; the tension it also computes is not stacked)."""
    l, vs, vw, th, ga, dth, dga = _valid_terms(df)
    return np.hstack([th * vs / (l + 1e-8), vs / (l + 1e-8), th * vs, vs, vs ** 2 / (l + 1e-8), th, ga,
                      ga * vw / (l + 1e-8), vw / (l + 1e-8), ga * vw, vw, dth, dga])


def build_gamma_features_valid(df, L=3.0, cable_wet_weight=1.521):
    """main_fun.py:574-642."""
    l, vs, vw, th, ga, dth, dga = _valid_terms(df)
    return np.hstack([ga * vw / (l + 1e-8), vw / (l + 1e-8), ga * vw, vw, vw ** 2 / (l + 1e-8), th, ga,
                      th * vs / (l + 1e-8), vs / (l + 1e-8), th * vs, vs, dth, dga])


def compute_derivatives(df):
    """main_fun.py:645-655: second derivatives of the Savitzky-Golay (11, 3) smoothed angles."""
    t = np.asarray(df["Time"], np.float64)
    out = []
    for k in ("Theta", "Gamma"):
        s = savgol_filter(np.asarray(df[k], np.float64), 11, 3)
        out.append(np.gradient(np.gradient(s, t), t))
    return out[0], out[1]


def preprocess_signals(df, sigma=2):
    """main_fun.py:768-776."""
    return (np.asarray(df["Time"], np.float64), gaussian_filter1d(np.asarray(df["Theta"], np.float64), sigma),
            gaussian_filter1d(np.asarray(df["Gamma"], np.float64), sigma))


def features_dd(df):
    """main_fun.py:811-871: (features (T,14), targets (T,2)); V1 in m/s here."""
    P0, P1, V1mm, _, t, th, ga = _experiment_arrays(df)
    V1 = V1mm / 1000
    a = np.stack([np.gradient(V1[:, j], t) for j in range(3)], 1)
    th_s, ga_s = savgol_filter(th, 11, 3), savgol_filter(ga, 11, 3)
    th_d, ga_d = np.gradient(th_s, t), np.gradient(ga_s, t)
    th_dd, ga_dd = np.gradient(th_d, t), np.gradient(ga_d, t)
    rel = P1 - P0
    unit = rel / (np.linalg.norm(rel, axis=1, keepdims=True) + 1e-8)
    v_surge = np.sum(V1 * unit, axis=1)
    v_sway = np.linalg.norm(np.cross(V1, unit), axis=1)
    a_surge, a_sway = np.gradient(v_surge, t), np.gradient(v_sway, t)
    col = lambda v: np.asarray(v).reshape(-1, 1)
    feats = np.hstack([col(th_s), col(ga_s), col(th_d), col(ga_d), col(v_sway), col(v_surge), col(a_sway), col(a_surge),
                       V1, a])
    return feats, np.hstack([col(th_dd), col(ga_dd)])
