"""Receding-horizon controller around the sampling-MPC step -- the role the reference's
absent ``pympc`` controller played at its call sites.  Plant state, nominal control sequence, step counter and
log live in HBM (cat_ctrl_*); a control step is 2 kernels (3 with MPPI) replayed as one CUDA graph, so a run of
n steps costs one upload, n graph launches and one read-back.  Definition: oracle ``mpc_receding_horizon``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .context import default_context
from .mpc import make_params, make_problems

LOG_COLUMNS = ("P1x", "P1y", "P1z", "ux", "uy", "uz", "theta", "gamma", "C", "z_low", "best_cost", "best_idx",
               "step_cost", "step", "span", "reserved")


class RecedingHorizonController:
    """``RecedingHorizonController(params, p_ref_traj, K, H, sigma, u_max)``; ``run(n)`` enqueues n closed-loop
    control steps (plant = the model), ``read()`` synchronises and returns the log / state / nominal sequence.

    params      dict as ``default_mpc_params()`` (initial pose, weights, dt, L); per-problem values may be
                (n,3)/(n,) arrays when ``n_problems`` > 1
    p_ref_traj  (ref_len,3) or (n_problems,ref_len,3) target positions in m; control step t tracks rows
                t .. t+H-1 (the last row is held)
    sigma,u_max (3,) mm/s: sampling spread around the nominal sequence, bound on the commanded velocity
    mode        'argmin' (next nominal = best candidate) | 'mppi' (exp(-(J-Jmin)/lam)-weighted mean)
    """

    def __init__(self, params, p_ref_traj, K, H, sigma, u_max, *, mode="argmin", lam=1.0, seed=0, clip=3.0,
                 max_steps=1024, u_nom0=None, problems=None, use_graph=True, dtype=np.float64, ctx=None):
        self.ctx = ctx or default_context()
        self.dtype = np.dtype(dtype)
        if self.dtype not in (np.dtype(np.float64), np.dtype(np.float32)):
            raise ValueError("dtype must be float64 or float32")
        if mode not in ("argmin", "mppi"):
            raise ValueError("mode must be 'argmin' or 'mppi'")
        ref = np.ascontiguousarray(p_ref_traj, np.float64)
        if ref.ndim == 2:
            ref = ref[None]
        if ref.ndim != 3 or ref.shape[2] != 3:
            raise ValueError("p_ref_traj must be (ref_len,3) or (n_problems,ref_len,3)")
        n, ref_len = ref.shape[0], ref.shape[1]
        probs = make_problems(params, n) if problems is None else np.ascontiguousarray(problems, _lib.PROBLEM_DTYPE)
        if probs.shape != (n,):
            raise ValueError("problems must have one entry per reference trajectory")
        self.K, self.H, self.n, self.max_steps = int(K), int(H), int(n), int(max_steps)
        self._params = make_params(params, K, H, n_problems=n, flags=_lib.MPC_RNG_CONTROLS, seed=seed, sigma=sigma,
                                   clip=clip)
        cfg = _lib.CatCtrlConfig()
        cfg.mode = _lib.CTRL_MPPI if mode == "mppi" else _lib.CTRL_ARGMIN
        cfg.max_steps = self.max_steps
        cfg.use_graph = 1 if use_graph else 0
        cfg.lam = float(lam)
        for a in range(3):
            cfg.u_max[a] = float(np.broadcast_to(np.asarray(u_max, np.float64), (3,))[a])
        un = None
        if u_nom0 is not None:
            un = np.ascontiguousarray(np.broadcast_to(np.asarray(u_nom0, np.float64), (n, self.H, 3)))
        h = C.c_void_p()
        fn = self.ctx.lib.cat_ctrl_create_f64 if self.dtype == np.float64 else self.ctx.lib.cat_ctrl_create_f32
        self.ctx.check(fn(self.ctx.handle, C.byref(self._params), C.byref(cfg), _lib.ptr(probs), _lib.ptr(un),
                          _lib.ptr(ref), ref_len, C.byref(h)))
        self.handle = h

    def run(self, n_steps):
        """Enqueue n_steps control steps on the context stream (no host synchronisation)."""
        self.ctx.check(self.ctx.lib.cat_ctrl_run(self.handle, int(n_steps)))
        return self

    def set_state(self, problems):
        """Overwrite the plant state with a measurement (PROBLEM_DTYPE array, one entry per problem)."""
        pr = np.ascontiguousarray(problems, _lib.PROBLEM_DTYPE)
        if pr.shape != (self.n,):
            raise ValueError("one state per problem")
        self.ctx.check(self.ctx.lib.cat_ctrl_set_state(self.handle, _lib.ptr(pr)))

    def read(self, log_from=0, log_n=None):
        """Synchronise; -> dict(step, state (PROBLEM_DTYPE (n,)), u_nom (n,H,3), log (n,log_n,16))."""
        step = C.c_int32()
        state = np.zeros(self.n, _lib.PROBLEM_DTYPE)
        u_nom = np.empty((self.n, self.H, 3))
        self.ctx.check(self.ctx.lib.cat_ctrl_read(self.handle, C.byref(step), _lib.ptr(state), _lib.ptr(u_nom), None,
                                                  0, 0))
        if log_n is None:
            log_n = max(0, min(int(step.value), self.max_steps) - int(log_from))
        log = np.empty((self.n, int(log_n), _lib.CTRL_LOG_W))
        if log_n:
            self.ctx.check(self.ctx.lib.cat_ctrl_read(self.handle, None, None, None, _lib.ptr(log), int(log_from),
                                                      int(log_n)))
        return dict(step=int(step.value), state=state, u_nom=u_nom, log=log)

    def close(self):
        if getattr(self, "handle", None) and self.ctx.handle:
            self.ctx.lib.cat_ctrl_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def mpc_receding_horizon(params, p_ref_traj, n_steps, K, H, sigma, u_max, *, seed=0, mode="argmin", lam=1.0,
                         clip=3.0, u_nom0=None, dtype=np.float64, use_graph=True, ctx=None):
    """Closed loop of n_steps control steps for one problem -> dict(log (n_steps,16), u_nom (H,3), state)."""
    c = RecedingHorizonController(params, p_ref_traj, K, H, sigma, u_max, mode=mode, lam=lam, seed=seed, clip=clip,
                                  max_steps=max(1, int(n_steps)), u_nom0=u_nom0, use_graph=use_graph, dtype=dtype,
                                  ctx=ctx)
    try:
        r = c.run(n_steps).read()
    finally:
        c.close()
    return dict(log=r["log"][0], u_nom=r["u_nom"][0], state=r["state"][0], step=r["step"])
