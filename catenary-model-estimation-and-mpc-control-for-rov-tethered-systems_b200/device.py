"""Device-resident sampling-MPC step: buffers live in HBM as torch tensors (torch is only the allocator and
the stream/NCCL plumbing); each ``launch`` is ONE kernel through cat_mpc_step_dev_* on torch's current stream.
Used by the receding-horizon loop, the multi-GPU shard exchange and bench.py's device-timed leg."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .context import Context
from .mpc import _wire_controls, make_params, make_problems


class MpcDeviceStep:
    def __init__(self, params, p_ref, K, *, controls=None, layout="H3K", sigma=None, seed=0, clip=3.0, k_offset=0,
                 problems=None, problem_offset=0, keep_costs=False, dtype=np.float64, device=0, ctx=None,
                 controls_dtype=None, private_workspace=False):
        import torch
        self.torch = torch
        self.dev = torch.device("cuda", device)
        self.ctx = ctx or Context(device)
        self.dtype = np.dtype(dtype)
        tdt = torch.float64 if self.dtype == np.float64 else torch.float32
        p_ref = np.ascontiguousarray(p_ref, np.float64)
        if p_ref.ndim == 2:
            p_ref = p_ref[None]
        n, H = p_ref.shape[0], p_ref.shape[1]
        probs = make_problems(params, n) if problems is None else np.ascontiguousarray(problems, _lib.PROBLEM_DTYPE)
        flags = 0
        if controls is None:
            if sigma is None:
                raise ValueError("controls=None needs sigma")
            flags |= _lib.MPC_RNG_CONTROLS
            self.controls = None
        else:
            if layout == "KH3":
                flags |= _lib.MPC_CONTROLS_AOS
            U, wire = _wire_controls(controls, self.dtype, controls_dtype)   # float32 storage under fp64 arithmetic
            flags |= wire
            self.controls = torch.as_tensor(U).to(self.dev)
        self.params = make_params(params, K, H, n_problems=n, k_offset=k_offset, problem_offset=problem_offset,
                                  flags=flags, seed=seed, sigma=(0, 0, 0) if sigma is None else sigma, clip=clip)
        self.K, self.H, self.n = int(K), int(H), int(n)
        self.problems = torch.from_numpy(probs.view(np.uint8).reshape(n, -1).copy()).to(self.dev)
        self.p_ref = torch.from_numpy(p_ref).to(self.dev)
        # winner = [cost bits (n) | idx (n)] in ONE int64 buffer, so a single 16-byte message per problem
        # carries it across ranks (distributed.exchange_winner)
        self.winner = torch.zeros(2 * n, dtype=torch.int64, device=self.dev)
        self.best_cost = self.winner[:n].view(torch.float64)
        self.best_idx = self.winner[n:]
        self.costs = torch.empty((n, K), dtype=tdt, device=self.dev) if keep_costs else None
        self._host_winner = None
        self._fn = self.ctx.lib.cat_mpc_step_dev_f64 if self.dtype == np.float64 else self.ctx.lib.cat_mpc_step_dev_f32
        # per-block winner slots + tickets of THIS step object: steps that share a Context may then be in flight on
        # different streams at once (the context's own scratch is shared by every launch on it)
        self._ws = None
        if private_workspace:
            h = C.c_void_p()
            self.ctx.check(self.ctx.lib.cat_mpc_workspace_create(self.ctx.handle, self.K, self.n, C.byref(h)))
            self._ws = h

    def set_controls(self, controls_dev):
        """Swap in another device tensor of candidate controls (same layout and dtype)."""
        self.controls = controls_dev

    def host_winner(self):
        """int64 [2n] view of page-locked, device-mapped host memory: with ``launch(to_host=True)`` the kernel's
        last block writes [cost bits | idx] straight into it over PCIe, so the step needs no D2H copy."""
        if self._host_winner is None:
            self._host_winner = self.ctx.pinned_empty((2 * self.n,), np.int64)
            self._host_winner[:] = 0
        return self._host_winner

    def launch(self, winner=None, to_host=False):
        """Enqueue the step on torch's current stream; returns immediately.  ``winner``: alternative int64
        [2n] device buffer for [cost bits | idx] (double-buffering across pipelined steps); ``to_host``: write
        the winner into ``host_winner()`` instead (read it with ``result_host``)."""
        torch = self.torch
        if to_host:
            wp = self.host_winner().ctypes.data
        else:
            wp = (self.winner if winner is None else winner).data_ptr()
        self.ctx.set_stream(torch.cuda.current_stream(self.dev).cuda_stream)
        self.ctx.lib.cat_set_workspace(self.ctx.handle, self._ws)
        self.ctx.check(self._fn(self.ctx.handle, C.byref(self.params), self.problems.data_ptr(),
                                None if self.controls is None else self.controls.data_ptr(), self.p_ref.data_ptr(),
                                wp, wp + 8 * self.n,
                                None if self.costs is None else self.costs.data_ptr()))

    def result_host(self):
        """(best_idx (n,), best_cost (n,)) of the last ``launch(to_host=True)``: a stream synchronise, no copy."""
        self.torch.cuda.current_stream(self.dev).synchronize()
        w = self.host_winner()
        return w[self.n:].copy(), w[:self.n].view(np.float64).copy()

    def result(self):
        """(best_idx (n,), best_cost (n,)) on the host (one small D2H + sync)."""
        w = self.winner.cpu().numpy()
        return w[self.n:].copy(), w[:self.n].view(np.float64).copy()

    def close(self):
        if self._ws is not None and self.ctx.handle:
            self.ctx.lib.cat_set_workspace(self.ctx.handle, None)
            self.ctx.lib.cat_mpc_workspace_destroy(self.ctx.handle, self._ws)
        self._ws = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def sample_controls(params, K, H, sigma, *, seed=0, clip=3.0, k_offset=0, problem_offset=0, n_problems=1, device=0,
                    ctx=None, to_host=True):
    """The candidate controls the rollout kernel draws on the device, materialised (cat_sample_controls_dev_f64):
    sigma * clip(N(0,1), +-clip) from Philox4x32-10 keyed by (seed, problem, global k, h), mm/s.  Returns
    (n_problems, H, 3, K) float64 -- the kernel's SoA layout -- as a NumPy array, or as the device tensor with
    ``to_host=False``.  The definition is oracle.sample_controls."""
    import torch
    dev = torch.device("cuda", device)
    ctx = ctx or Context(device)
    prm = make_params(params, K, H, n_problems=n_problems, k_offset=k_offset, problem_offset=problem_offset,
                      flags=_lib.MPC_RNG_CONTROLS, seed=seed, sigma=sigma, clip=clip)
    out = torch.empty((int(n_problems), int(H), 3, int(K)), dtype=torch.float64, device=dev)
    ctx.set_stream(torch.cuda.current_stream(dev).cuda_stream)
    ctx.check(ctx.lib.cat_sample_controls_dev_f64(ctx.handle, C.byref(prm), out.data_ptr()))
    return out.cpu().numpy() if to_host else out
