"""Sampling-MPC step over a batch of candidate control sequences (new composition).

The definition of the rollout, the cost and the argmin is oracle/catenary_oracle.py::mpc_rollout_costs; the
CUDA kernel (csrc/cat_kernels_mpc.cuh) follows it.  NumPy in / NumPy out here; ``MpcDeviceStep`` keeps the
buffers on the device for the receding-horizon / multi-GPU callers.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .context import default_context
from .dynamics import equation_id

C_LO, C_HI = 1e-6, 10.0      # main_fun.py:425
NAN_PENALTY = 1e12


def default_mpc_params(**over):
    """Config-3 constants; pose of fully_augmented_catenary.py:11-15."""
    p = dict(dt=0.01, L=3.0, P0=(0.0, 0.0, 0.0), P1=(1.0, 1.0, -1.0), V1=(0.0, 0.0, 0.0), A1=(0.0, 0.0, 0.0),
             theta=float(np.radians(-10.0)), gamma=float(np.radians(20.0)),
             theta_prev=float(np.radians(-10.0)), gamma_prev=float(np.radians(20.0)),
             w_trk=1.0, w_du=1e-4, w_len=1e3, w_sag=1e3, w_ang=0.1, rho=0.95, z_floor=-2.0, margin=0.1,
             c_lo=C_LO, c_hi=C_HI, nan_penalty=NAN_PENALTY, theta_cx=None, gamma_cx=None,
             theta_expr=None, gamma_expr=None)
    p.update(over)
    return p


def circle_reference(H, dt, center=(1.0, 1.0, -1.0), radius=0.1, period=10.0):
    """Target P_ref[h], h = 1..H: 'case 12' circle of Rov_traj_gen.py:113-117 through the initial pose."""
    t = (np.arange(H) + 1) * dt
    c = np.asarray(center, np.float64)
    w = 2 * np.pi / period
    return np.stack([c[0] - radius + radius * np.cos(w * t), c[1] + radius * np.sin(w * t),
                     np.full(H, c[2])], axis=1)


def make_params(params, K, H, *, n_problems=1, k_offset=0, problem_offset=0, flags=0, seed=0,
                sigma=(0.0, 0.0, 0.0), clip=3.0):
    p = params
    s = _lib.CatMpcParams()
    s.K, s.k_offset, s.H, s.n_problems = int(K), int(k_offset), int(H), int(n_problems)
    s.flags, s.problem_offset = int(flags), int(problem_offset)
    s.eq_theta = -1 if p.get("theta_cx") is None else equation_id("dtheta_dt", p["theta_cx"])
    s.eq_gamma = -1 if p.get("gamma_cx") is None else equation_id("dgamma_dt", p["gamma_cx"])
    for k in ("dt", "L", "w_trk", "w_du", "w_len", "w_sag", "w_ang", "rho", "z_floor", "margin", "c_lo", "c_hi",
              "nan_penalty"):
        setattr(s, k, float(p[k]))
    s.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    s.sigma[0], s.sigma[1], s.sigma[2] = (float(v) for v in sigma)
    s.clip = float(clip)
    return s


def make_problems(params, n_problems=1):
    """Structured array (PROBLEM_DTYPE) from a params dict; per-problem values may be given as (n,3)/(n,)."""
    # the record is 16 packed doubles (P0 P1 V1 A1 theta gamma theta_prev gamma_prev): filled through a (n,16) view,
    # a third of the cost of per-field structured assignment (this runs on every host MPC step)
    flat = np.empty((n_problems, 16))
    for j, k in enumerate(("P0", "P1", "V1", "A1")):
        flat[:, 3 * j:3 * j + 3] = params[k]
    for j, k in enumerate(("theta", "gamma", "theta_prev", "gamma_prev")):
        flat[:, 12 + j] = params[k]
    return flat.view(_lib.PROBLEM_DTYPE).reshape(n_problems)


def _programs(params):
    """(theta Program, gamma Program) when params carries run-time equations ('theta_expr' / 'gamma_expr': an
    equation string over x0..x17, an expr.Program or an expr.ExpressionModel), else None."""
    te, ge = params.get("theta_expr"), params.get("gamma_expr")
    if te is None and ge is None:
        return None
    if te is None or ge is None:
        raise ValueError("theta_expr and gamma_expr go together")
    from .expr import Program, compile_expression

    def prog(e):
        e = getattr(e, "program", e)
        if not isinstance(e, Program):
            e = compile_expression(e, 18)
        if e.n_features != 18:
            raise ValueError("MPC equations are over the 18-feature layout of simple.py:15-41")
        return e
    return prog(te), prog(ge)


def _wire_controls(controls, dtype, controls_dtype):
    """(array handed to the library, extra flags).  The rollout arithmetic is ``dtype``; the candidates may be stored
    -- and cross PCIe -- as float32 while the arithmetic stays fp64 (``controls_dtype=np.float32``, or simply a
    float32 array): half the bytes of the PCIe-bound host step, and exact parity as long as the oracle is given the
    same rounded values."""
    dtype = np.dtype(dtype)
    if controls_dtype is None:
        controls_dtype = np.float32 if (dtype == np.float64 and getattr(controls, "dtype", None) == np.float32) else dtype
    controls_dtype = np.dtype(controls_dtype)
    if dtype == np.float32:
        controls_dtype = dtype
    if controls_dtype not in (np.dtype(np.float32), np.dtype(np.float64)):
        raise ValueError("controls_dtype must be float32 or float64")
    U = np.ascontiguousarray(controls, controls_dtype)
    return U, (_lib.MPC_CONTROLS_F32 if (dtype == np.float64 and controls_dtype == np.float32) else 0)


def mpc_sample_step(params, controls, p_ref, *, layout="KH3", return_costs=True, dtype=np.float64, seed=0,
                    sigma=None, clip=3.0, K=None, k_offset=0, ctx=None, controls_dtype=None, p2p=None):
    """One sampling-MPC step -> (best_idx, best_cost, costs | None).

    controls: (K,H,3) [layout 'KH3', the oracle's] or (H,3,K) [layout 'H3K', the kernel's coalesced SoA]
    commanded V1 in mm/s; or None with ``sigma`` (3,) and ``K`` to sample on the device (Philox4x32-10 and the
    tabulated inverse normal CDF, keyed by (seed, problem, k, h)).  p_ref (H,3) in m.  best_idx: lowest cost, ties
    -> lowest index, NaN costs rank last.  ``controls_dtype=np.float32`` (or a float32 array) stores the candidates
    as float under fp64 arithmetic.  ``p2p``: a distributed.P2PWinnerExchange -- this rank's shard goes in, the
    JOB-WIDE winner comes out on every rank (exchanged inside the kernel over NVLink)."""
    ctx = ctx or default_context()
    dtype = np.dtype(dtype)
    p_ref = np.ascontiguousarray(p_ref, np.float64)
    H = p_ref.shape[0]
    flags = 0
    if controls is None:
        if sigma is None or K is None:
            raise ValueError("controls=None needs sigma and K")
        flags |= _lib.MPC_RNG_CONTROLS
        U = None
    else:
        U, wire = _wire_controls(controls, dtype, controls_dtype)
        flags |= wire
        if U is controls:                                # the caller's own array (no conversion copy): page-lock it once
            ctx.maybe_register(U)
        if layout == "KH3":
            if U.ndim != 3 or U.shape[1:] != (H, 3):
                raise ValueError(f"controls must be (K,{H},3)")
            K = U.shape[0]
            flags |= _lib.MPC_CONTROLS_AOS
        elif layout == "H3K":
            if U.ndim != 3 or U.shape[:2] != (H, 3):
                raise ValueError(f"controls must be ({H},3,K)")
            K = U.shape[2]
        else:
            raise ValueError("layout must be 'KH3' or 'H3K'")
    s = make_params(params, K, H, k_offset=k_offset, flags=flags, seed=seed, sigma=(0, 0, 0) if sigma is None else sigma, clip=clip)
    pr = make_problems(params)
    best_cost = np.empty(1)
    best_idx = np.empty(1, np.int64)
    costs = np.empty(K, dtype) if return_costs else None
    progs = _programs(params)
    if progs is not None and p2p is not None:
        raise ValueError("run-time equation programs and the fused exchange are not combined in one call")
    if progs is not None:
        # run-time equation programs (a retrained model's equation strings) instead of the generated code
        pt, pg = progs
        fn = ctx.lib.cat_mpc_step_expr_f64 if dtype == np.float64 else ctx.lib.cat_mpc_step_expr_f32
        ctx.check(fn(ctx.handle, C.byref(s), _lib.ptr(pr), _lib.ptr(U), _lib.ptr(p_ref),
                     len(pt.ops), _lib.ptr(pt.ops), len(pt.consts), _lib.ptr(pt.consts) if len(pt.consts) else None,
                     len(pg.ops), _lib.ptr(pg.ops), len(pg.consts), _lib.ptr(pg.consts) if len(pg.consts) else None,
                     _lib.ptr(best_cost), _lib.ptr(best_idx), _lib.ptr(costs)))
        return int(best_idx[0]), float(best_cost[0]), costs
    if p2p is not None:
        fn = ctx.lib.cat_mpc_step_p2p_f64 if dtype == np.float64 else ctx.lib.cat_mpc_step_p2p_f32
        ctx.check(fn(ctx.handle, p2p.handle, C.byref(s), _lib.ptr(pr), _lib.ptr(U), _lib.ptr(p_ref),
                     _lib.ptr(best_cost), _lib.ptr(best_idx), _lib.ptr(costs)))
        p2p.check_winner(int(best_idx[0]))
        return int(best_idx[0]), float(best_cost[0]), costs
    fn = ctx.lib.cat_mpc_step_f64 if dtype == np.float64 else ctx.lib.cat_mpc_step_f32
    ctx.check(fn(ctx.handle, C.byref(s), _lib.ptr(pr), _lib.ptr(U), _lib.ptr(p_ref), _lib.ptr(best_cost),
                 _lib.ptr(best_idx), _lib.ptr(costs)))
    return int(best_idx[0]), float(best_cost[0]), costs


def mpc_fleet_step(params, problems, p_ref, K, sigma, *, seed=0, clip=3.0, problem_offset=0, return_costs=False,
                   dtype=np.float64, ctx=None):
    """Independent MPC problems in one launch (config 5): problems = PROBLEM_DTYPE array (n,), p_ref (n,H,3),
    controls sampled on the device.  -> (best_idx (n,), best_cost (n,), costs (n,K) | None)."""
    ctx = ctx or default_context()
    dtype = np.dtype(dtype)
    problems = np.ascontiguousarray(problems, _lib.PROBLEM_DTYPE)
    n = problems.shape[0]
    p_ref = np.ascontiguousarray(p_ref, np.float64)
    if p_ref.ndim != 3 or p_ref.shape[0] != n or p_ref.shape[2] != 3:
        raise ValueError("p_ref must be (n_problems,H,3)")
    H = p_ref.shape[1]
    s = make_params(params, K, H, n_problems=n, problem_offset=problem_offset, flags=_lib.MPC_RNG_CONTROLS,
                    seed=seed, sigma=sigma, clip=clip)
    best_cost = np.empty(n)
    best_idx = np.empty(n, np.int64)
    costs = np.empty((n, K), dtype) if return_costs else None
    fn = ctx.lib.cat_mpc_step_f64 if dtype == np.float64 else ctx.lib.cat_mpc_step_f32
    ctx.check(fn(ctx.handle, C.byref(s), _lib.ptr(problems), None, _lib.ptr(p_ref), _lib.ptr(best_cost),
                 _lib.ptr(best_idx), _lib.ptr(costs)))
    return best_idx, best_cost, costs


def reduce_winners(costs, idxs):
    """Deterministic fold of per-shard winners (lowest cost, ties -> lowest global index, NaN ranks last)."""
    lib = _lib.load()
    c = np.ascontiguousarray(costs, np.float64)
    i = np.ascontiguousarray(idxs, np.int64)
    bc = np.empty(1)
    bi = np.empty(1, np.int64)
    rc = lib.cat_reduce_winners_f64(c.size, _lib.ptr(c), _lib.ptr(i), _lib.ptr(bc), _lib.ptr(bi))
    if rc != 0:
        raise ValueError("reduce_winners: empty input")
    return int(bi[0]), float(bc[0])
