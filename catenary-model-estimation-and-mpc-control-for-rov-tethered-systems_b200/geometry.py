"""Reference call surface for the catenary geometry, backed by the CUDA library.

Same names / argument meaning / return shapes as the reference:
  Catenary(length, reference_frame)(a, b)         pympc.models.catenary (call sites catenary.py:10,25,51)
  rodrigues_rotation(vector, axis, angle_rad)     main_fun.py:18-35
  transform_catenary(A, B, catenary_fn, th, ga)   main_fun.py:38-111
  solve_catenary(l, delta_H, L)                   main_fun.py:418-431
plus batched forms (leading batch dimension) that the reference's per-element Python loops lack.
"""
from __future__ import annotations

import numpy as np

from . import _lib
from .context import default_context

N_PTS = 100     # catenary_model.py:28, catenary_3d.py:52


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None and a.shape != shape:
        raise ValueError(f"expected shape {shape}, got {a.shape}")
    return a


def solve_catenary(l, delta_H, L, *, dtype=np.float64, ctx=None):
    """main_fun.py:418-431: root C in (1e-6, 10) of C^2 (L^2 - dH^2) - 4 sinh^2(l C / 2); nan on failure.
    Array or scalar l / delta_H; L scalar or array.  Returns an ndarray of l's shape (0-d for scalars,
    as np.vectorize does)."""
    ctx = ctx or default_context()
    dtype = np.dtype(dtype)
    if dtype not in (np.dtype(np.float64), np.dtype(np.float32)):
        raise ValueError("dtype must be float64 or float32")
    l_a, h_a = np.broadcast_arrays(np.asarray(l, dtype), np.asarray(delta_H, dtype))
    shape = l_a.shape
    l_c = np.ascontiguousarray(l_a).ravel()
    h_c = np.ascontiguousarray(h_a).ravel()
    L_a = np.asarray(L, dtype)
    if L_a.ndim == 0:
        L_c, stride = L_a.reshape(1).copy(), 0
    else:
        L_c, stride = np.ascontiguousarray(np.broadcast_to(L_a, shape)).ravel(), 1
    out = np.empty(l_c.shape, dtype)
    fn = ctx.lib.cat_solve_param_f64 if dtype == np.float64 else ctx.lib.cat_solve_param_f32
    ctx.check(fn(ctx.handle, l_c.size, _lib.ptr(l_c), _lib.ptr(h_c), _lib.ptr(L_c), stride, _lib.ptr(out)))
    return out.reshape(shape)


def span_terms(l, delta_H, L, w=1.521, *, ctx=None):
    """(C, x_low, z_low_rel, tension) per span.  tension: main_fun.py:302-305 (w: test_cluster.py:23)."""
    ctx = ctx or default_context()
    l_a, h_a = np.broadcast_arrays(np.asarray(l, np.float64), np.asarray(delta_H, np.float64))
    shape = l_a.shape
    l_c, h_c = np.ascontiguousarray(l_a).ravel(), np.ascontiguousarray(h_a).ravel()
    L_a = np.asarray(L, np.float64)
    if L_a.ndim == 0:
        L_c, stride = L_a.reshape(1).copy(), 0
    else:
        L_c, stride = np.ascontiguousarray(np.broadcast_to(L_a, shape)).ravel(), 1
    outs = [np.empty(l_c.size) for _ in range(4)]
    ctx.check(ctx.lib.cat_span_terms_f64(ctx.handle, l_c.size, _lib.ptr(l_c), _lib.ptr(h_c), _lib.ptr(L_c), stride,
                                         float(w), *[_lib.ptr(o) for o in outs]))
    return tuple(o.reshape(shape) for o in outs)


def rodrigues_rotation(vector, axis, angle_rad, *, ctx=None):
    """main_fun.py:18-35.  vector (3,) or (n,3); axis (3,) or (n,3); angle scalar or (n,)."""
    ctx = ctx or default_context()
    v = np.asarray(vector, np.float64)
    single = v.ndim == 1
    v2 = _f64(v.reshape(-1, 3))
    n = v2.shape[0]
    ax = np.asarray(axis, np.float64)
    if ax.ndim == 1:
        ax_c, a_stride = _f64(ax, (3,)), 0
    else:
        ax_c, a_stride = _f64(ax, (n, 3)), 1
    an = np.asarray(angle_rad, np.float64)
    if an.ndim == 0:
        an_c, g_stride = an.reshape(1).copy(), 0
    else:
        an_c, g_stride = _f64(an, (n,)), 1
    out = np.empty((n, 3))
    ctx.check(ctx.lib.cat_rodrigues_f64(ctx.handle, n, _lib.ptr(v2), _lib.ptr(ax_c), a_stride, _lib.ptr(an_c),
                                        g_stride, None, _lib.ptr(out)))
    return out[0] if single else out


def catenary_shapes(A, B, L, theta=None, gamma=None, *, n_pts=N_PTS, which=_lib.OUT_FINAL, dtype=np.float64,
                    ctx=None, out=None):
    """Batched Catenary.__call__ + transform_catenary: A, B (n,3); L scalar or (n,); theta, gamma (n,) or None.

    Returns (curves, C, valid): curves is a list with one (n, n_pts, 3) array per bit of ``which``
    (order original, theta_rotated, theta_aligned, final), C (n,), valid (n,) uint8 (bit 0: span A->B in the
    solve domain, bit 1: rotated span).  ``out``: a caller-owned (n_outputs, n, n_pts, 3) array to fill instead of a
    fresh one -- for a million curves the 2.4 GB result is the cost of the call, and a reused page-locked buffer
    (``ctx.pinned_empty``) takes it at PCIe speed where a fresh NumPy array pays page faults and the pageable path."""
    ctx = ctx or default_context()
    dtype = np.dtype(dtype)
    A_c = np.ascontiguousarray(A, dtype).reshape(-1, 3)
    B_c = np.ascontiguousarray(B, dtype).reshape(-1, 3)
    n = A_c.shape[0]
    if B_c.shape[0] != n:
        raise ValueError("A and B must have the same number of rows")
    L_a = np.asarray(L, dtype)
    if L_a.ndim == 0:
        L_c, stride = L_a.reshape(1).copy(), 0
    else:
        L_c, stride = np.ascontiguousarray(L_a).reshape(n), 1
    th = None if theta is None else np.ascontiguousarray(np.broadcast_to(np.asarray(theta, dtype), (n,)))
    ga = None if gamma is None else np.ascontiguousarray(np.broadcast_to(np.asarray(gamma, dtype), (n,)))
    nout = bin(which & 15).count("1")
    if nout == 0 or which != (which & 15):
        raise ValueError("which must be a non-empty mask of OUT_* bits")
    if out is None:
        out = np.empty((nout, n, n_pts, 3), dtype)
    elif out.shape != (nout, n, n_pts, 3) or out.dtype != dtype or not out.flags["C_CONTIGUOUS"]:
        raise ValueError(f"out must be a C-contiguous {dtype} array of shape {(nout, n, n_pts, 3)}")
    C_out = np.empty(n, dtype)
    valid = np.empty(n, np.uint8)
    fn = ctx.lib.cat_shape_f64 if dtype == np.float64 else ctx.lib.cat_shape_f32
    ctx.check(fn(ctx.handle, n, _lib.ptr(A_c), _lib.ptr(B_c), _lib.ptr(L_c), stride, _lib.ptr(th), _lib.ptr(ga),
                 int(n_pts), int(which), _lib.ptr(out), _lib.ptr(C_out), _lib.ptr(valid)))
    return [out[i] for i in range(nout)], C_out, valid


def lowest_point(curves, *, ctx=None):
    """np.argmin(curve[:,2]) (fully_augmented_catenary.py:20-22) for (n_pts,3) or (n,n_pts,3) curves.
    Returns (index, point)."""
    ctx = ctx or default_context()
    c = np.asarray(curves, np.float64)
    single = c.ndim == 2
    c3 = _f64(c.reshape((-1,) + c.shape[-2:]))
    n, n_pts, _ = c3.shape
    idx = np.empty(n, np.int64)
    pt = np.empty((n, 3))
    ctx.check(ctx.lib.cat_lowest_point_f64(ctx.handle, n, n_pts, _lib.ptr(c3), _lib.ptr(idx), _lib.ptr(pt)))
    return (int(idx[0]), pt[0]) if single else (idx, pt)


class Catenary:
    """Call surface of pympc.models.catenary.Catenary (catenary.py:10,25,51; source absent upstream, so the
    sampling is the builder's definition -- see DESIGN.md).  ``Catenary(3., 'ENU')(a, b)`` returns
    ``(C, H, D, points)``; points is (n_points,3) or None outside the solve domain (callers then use
    [a, b], main_fun.py:67-69)."""

    def __init__(self, length=3.0, reference_frame="ENU", n_points=N_PTS, ctx=None):
        if reference_frame not in ("ENU", "NED"):
            raise ValueError("reference_frame must be 'ENU' or 'NED'")
        self.length = float(length)
        self.reference_frame = reference_frame
        self.n_points = int(n_points)
        self._ctx = ctx

    def __call__(self, a, b):
        ctx = self._ctx or default_context()
        sgn = np.array([1.0, 1.0, 1.0 if self.reference_frame == "ENU" else -1.0])
        a2 = _f64(a, (3,)) * sgn
        b2 = _f64(b, (3,)) * sgn
        (pts,), C, valid = catenary_shapes(a2[None], b2[None], self.length, n_pts=self.n_points,
                                           which=_lib.OUT_ORIGINAL, ctx=ctx)
        if not (valid[0] & 1):
            return None, None, None, None
        rel = b2 - a2
        l = float(np.hypot(rel[0], rel[1]))
        _, x_low, z_low, _ = span_terms(l, rel[2], self.length, ctx=ctx)
        H = float(min(0.0, rel[2]) - z_low)
        D = float(min(max(float(x_low), 0.0), l))
        return float(C[0]), H, D, pts[0] * sgn


def transform_catenary(point_A, point_B, catenary_fn, theta_rad, gamma_rad, *, ctx=None):
    """main_fun.py:38-111 -> (original, theta_rotated, theta_aligned, final).

    With our ``Catenary`` as catenary_fn everything is one fused kernel; with any other callable the two
    curves come from it (None -> [start, end], main_fun.py:67-69) and the per-point rotations
    (main_fun.py:96-99,106-109) run on the device."""
    ctx = ctx or default_context()
    A = _f64(point_A, (3,))
    B = _f64(point_B, (3,))
    if isinstance(catenary_fn, Catenary) and catenary_fn.reference_frame == "ENU":
        curves, _, valid = catenary_shapes(A[None], B[None], catenary_fn.length, [theta_rad], [gamma_rad],
                                           n_pts=catenary_fn.n_points, which=15, ctx=ctx)
        orig, rot, ali, fin = (c[0] for c in curves)
        v = int(valid[0])
        if not (v & 1):
            orig = np.array([A, B])
        if not (v & 2):
            # straight-chord fallback has 2 points in the reference: rebuild the derived curves from its ends
            rot, ali, fin = rot[[0, -1]], ali[[0, -1]], fin[[0, -1]]
        return orig, rot, ali, fin

    def compute(start, end):
        res = catenary_fn(start, end)
        return np.asarray(res[3], np.float64) if res[3] is not None else np.array([start, end])

    original = compute(A, B)
    conn = B - A
    xy = np.array([conn[0], conn[1], 0.0])
    nxy = np.linalg.norm(xy)
    xy = np.array([1.0, 0.0, 0.0]) if nxy < 1e-9 else xy / nxy
    k_theta = np.cross(xy, [0.0, 0.0, 1.0])
    nk = np.linalg.norm(k_theta)
    k_theta = np.array([0.0, 1.0, 0.0]) if nk < 1e-9 else k_theta / nk
    rotated_B = A + rodrigues_rotation(conn, k_theta, theta_rad, ctx=ctx)
    theta_rotated = compute(A, rotated_B)
    aligned = np.empty_like(theta_rotated)
    final = np.empty_like(theta_rotated)
    n = theta_rotated.shape[0]
    ang = np.array([-float(theta_rad)])
    ctx.check(ctx.lib.cat_rodrigues_f64(ctx.handle, n, _lib.ptr(_f64(theta_rotated)), _lib.ptr(k_theta), 0,
                                        _lib.ptr(ang), 0, _lib.ptr(A), _lib.ptr(aligned)))
    k_gamma = conn / np.linalg.norm(conn)
    ang2 = np.array([float(gamma_rad)])
    ctx.check(ctx.lib.cat_rodrigues_f64(ctx.handle, n, _lib.ptr(aligned), _lib.ptr(k_gamma), 0, _lib.ptr(ang2), 0,
                                        _lib.ptr(A), _lib.ptr(final)))
    return original, theta_rotated, aligned, final


def compute_catenary_3D(p0, p1, rope_length, num_points, *, ctx=None):
    """models/catenary_3d.py:5-39: (3,),(3,),scalar,int -> (num_points,3); batched: (n,3),(n,3),(n,)|scalar ->
    (n,num_points,3).  Taut rope -> np.linspace(p0, p1); else the reference's fixed-point in ``a`` and symmetric
    sag profile."""
    ctx = ctx or default_context()
    a = np.ascontiguousarray(p0, np.float64)
    b = np.ascontiguousarray(p1, np.float64)
    single = a.ndim == 1
    a2, b2 = np.ascontiguousarray(a.reshape(-1, 3)), np.ascontiguousarray(b.reshape(-1, 3))
    if a2.shape != b2.shape:
        raise ValueError("p0 and p1 must have the same shape (..., 3)")
    n = a2.shape[0]
    L = np.ascontiguousarray(rope_length, np.float64).reshape(-1)
    if L.size not in (1, n):
        raise ValueError("rope_length must be a scalar or one value per rope")
    out = np.empty((n, int(num_points), 3))
    ctx.check(ctx.lib.cat_catenary3d_f64(ctx.handle, n, _lib.ptr(a2), _lib.ptr(b2), _lib.ptr(L),
                                         0 if L.size == 1 else 1, int(num_points), _lib.ptr(out)))
    return out[0] if single else out


def catenary_equation(x, C, *, ctx=None):
    """catenary_model.py:10-12: (cosh(C x) - 1) / C, elementwise (C scalar or like x)."""
    ctx = ctx or default_context()
    xa = np.ascontiguousarray(x, np.float64)
    Ca = np.ascontiguousarray(C, np.float64).reshape(-1)
    if Ca.size not in (1, xa.size):
        raise ValueError("C must be a scalar or have the shape of x")
    y = np.empty(xa.size)
    ctx.check(ctx.lib.cat_catenary_equation_f64(ctx.handle, xa.size, _lib.ptr(xa.reshape(-1)), _lib.ptr(Ca),
                                                0 if Ca.size == 1 else 1, _lib.ptr(y)))
    return y.reshape(xa.shape)


def solve_catenary_params(L, l, H, *, ctx=None):
    """catenary_model.py:14-20: the C with 2 sinh(C l / 2) / C = L on (1e-5, 10) -- the level-support case of
    ``solve_catenary`` (H is unused by the reference too).  The reference gets it from a bounded scalar
    MINIMISER of the squared residual with xatol 1e-5, so it agrees with this root to ~1e-5, not 1e-9; returns
    None where no root exists in the bracket (the reference's minimiser returns a bound-hugging value there)."""
    C = solve_catenary(l, 0.0, L, ctx=ctx)
    C = float(np.asarray(C).reshape(-1)[0]) if np.ndim(l) == 0 else C
    if np.ndim(l) == 0:
        return None if (np.isnan(C) or C < 1e-5) else C
    return C


def generate_catenary(x_range, L, l, H, *, ctx=None):
    """catenary_model.py:22-31: 100 samples of the level-support profile on [-l/2, l/2] -> (x, y)."""
    C = solve_catenary_params(L, l, H, ctx=ctx)
    if C is None:
        raise ValueError("Failed to compute catenary parameter.")
    x = np.linspace(-l / 2, l / 2, 100)
    return x, catenary_equation(x, C, ctx=ctx)
