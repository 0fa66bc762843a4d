"""Multi-GPU sharding of the sampling-MPC step: one process per GPU (torch.distributed), candidates split into
contiguous shards, ONE (cost, index) exchange per step.

NCCL has no MINLOC and an fp64 cost + int64 index do not pack into one 64-bit min, so every rank all-gathers
the 16-byte winners and folds them in rank order with the lowest-global-index tie-break
(cat_reduce_winners_f64; a NaN cost ranks with +inf) -- identical on every rank and identical to np.argmin over the concatenated costs.
The fleet configuration (independent problems) shards problems and needs no communication.
"""
from __future__ import annotations

import numpy as np

from . import _lib
from .mpc import reduce_winners

_P2P_LOCKSTEP, _P2P_PIPELINED = _lib.P2P_LOCKSTEP, _lib.P2P_PIPELINED


def shard_range(total, rank, world):
    """Contiguous shard [lo, hi) of ``total`` items for ``rank`` of ``world`` (remainder to the low ranks)."""
    base, rem = divmod(int(total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def exchange_winner(winner, group=None):
    """winner: int64 tensor [cost bits, global idx] of this rank (CPU tensor under gloo, CUDA under NCCL).
    Returns (best_idx, best_cost) of the whole job, the same on every rank."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    gathered = torch.empty(world * 2, dtype=torch.int64, device=winner.device)
    dist.all_gather_into_tensor(gathered, winner.contiguous(), group=group)
    g = gathered.cpu().numpy().reshape(world, 2)
    return reduce_winners(g[:, 0].copy().view(np.float64), g[:, 1].copy())


def pack_winner(cost, idx):
    """Host-side (cost, idx) -> int64[2] tensor (CPU); used by the gloo tests and by callers that already hold
    the local winner on the host."""
    import torch
    a = np.empty(2, np.int64)
    a[:1].view(np.float64)[0] = cost
    a[1] = idx
    return torch.from_numpy(a)


class ShardedMpcStep:
    """Config 4: K_total candidates of ONE problem split across the ranks of the default process group,
    controls sampled on the device from (seed, global k, h) so no scatter is needed."""

    def __init__(self, params, p_ref, K_total, sigma, *, seed=0, clip=3.0, dtype=np.float64, device=None,
                 group=None):
        import torch
        import torch.distributed as dist
        from .device import MpcDeviceStep
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        lo, hi = shard_range(K_total, self.rank, self.world)
        self.lo, self.hi = lo, hi
        device = torch.cuda.current_device() if device is None else device
        self.step = MpcDeviceStep(params, p_ref, hi - lo, sigma=sigma, seed=seed, clip=clip, k_offset=lo,
                                  dtype=dtype, device=device)

    def launch(self):
        self.step.launch()

    def finish(self):
        """Exchange the local winners; returns (best_idx, best_cost) of the whole job."""
        return exchange_winner(self.step.winner, self.group)

    def run(self):
        self.launch()
        return self.finish()


def fleet_shard(n_problems, rank, world):
    """Config 5: problems [lo, hi) of this rank; no cross-GPU communication."""
    return shard_range(n_problems, rank, world)


class PipelinedExchange:
    """Sync-free multi-GPU step: the MPC kernel runs on torch's current stream; the 16-byte all-gather and the
    device-side fold (cat_reduce_winners_dev_f64) run on a side stream, double-buffered, so step i+1's rollout
    overlaps step i's exchange and the host never blocks until ``result``.  Every rank ends with the same
    (cost bits, idx) pair in ``final[slot]``."""

    def __init__(self, step, group=None, depth=2):
        import torch
        import torch.distributed as dist
        from .context import Context
        self.torch, self.dist, self.step, self.group = torch, dist, step, group
        self.world = dist.get_world_size(group)
        dev = step.dev
        self.depth = depth
        self.side = torch.cuda.Stream(device=dev)
        self.fold_ctx = Context(dev.index)                  # its own handle: its stream is the side stream
        self.fold_ctx.set_stream(self.side.cuda_stream)
        self.local = [torch.zeros(2, dtype=torch.int64, device=dev) for _ in range(depth)]
        self.gathered = [torch.zeros(2 * self.world, dtype=torch.int64, device=dev) for _ in range(depth)]
        self.final = [torch.zeros(2, dtype=torch.int64, device=dev) for _ in range(depth)]
        self.done = [torch.cuda.Event() for _ in range(depth)]
        self.ready = [torch.cuda.Event() for _ in range(depth)]
        self.count = 0

    def submit(self):
        """Enqueue one step (rollout + exchange + fold); returns the slot holding its result."""
        torch = self.torch
        s = self.count % self.depth
        main = torch.cuda.current_stream(self.step.dev)
        if self.count >= self.depth:
            main.wait_event(self.done[s])                   # the exchange that last used this slot has drained
        self.step.launch(winner=self.local[s])
        self.ready[s].record(main)
        with torch.cuda.stream(self.side):
            self.side.wait_event(self.ready[s])
            self.dist.all_gather_into_tensor(self.gathered[s], self.local[s], group=self.group)
            self.fold_ctx.check(self.fold_ctx.lib.cat_reduce_winners_dev_f64(
                self.fold_ctx.handle, self.world, self.gathered[s].data_ptr(), self.final[s].data_ptr()))
            self.done[s].record(self.side)
        self.count += 1
        return s

    def result(self, slot):
        """(best_idx, best_cost) of the step submitted into ``slot`` (blocks the host on that slot only)."""
        self.done[slot].synchronize()
        w = self.final[slot].cpu().numpy()
        return int(w[1]), float(w[:1].view(np.float64)[0])

    def drain(self):
        self.side.synchronize()


class P2PWinnerExchange:
    """Fused multi-GPU step: ONE kernel per step per rank.  The rollout's last block posts this
    rank's (cost, global idx) into every peer's slot buffer over NVLink (CUDA IPC mappings) and folds the G
    winners in rank order, with no NCCL call and no second launch:

    * ``launch()`` (lock-step): the kernel waits for the peers' winners of the SAME step -- when it ends,
      ``step.winner`` holds the JOB-WIDE winner on every rank (same value as ``exchange_winner``).  The latency path.
    * ``launch(pipelined=True)``: the kernel folds the winners of the PREVIOUS step (posted a whole rollout ago, so
      nothing spins and back-to-back steps of G processes do not pay each other's launch jitter); the winner of
      step s arrives with step s+1, or with ``flush()`` after the last one.  The throughput path.
    * ``mpc.mpc_sample_step(..., p2p=self)``: NumPy in, job-wide winner out (cat_mpc_step_p2p_*, lock-step).

    A peer that does not show up within a few seconds makes the step return the sentinel (NaN, -1); ``result`` /
    ``result_host`` raise on it.  The process group is used once, at construction, to swap the 64-byte IPC handles.
    Every rank must launch the same sequence of steps and modes.  Not capturable into a CUDA graph."""

    def __init__(self, step=None, group=None, ctx=None, device=None):
        """``step``: the MpcDeviceStep whose launches carry the exchange (``launch``); for the host-pointer step (``mpc_sample_step(p2p=...)``) alone,
        a ``ctx`` (and ``device``) is enough."""
        import ctypes as C

        import torch
        import torch.distributed as dist
        self.step, self.torch = step, torch
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if step is not None and step.n != 1:
            raise ValueError("the fused exchange handles one problem per launch")
        ctx = step.ctx if step is not None else ctx
        if ctx is None:
            raise ValueError("P2PWinnerExchange needs a step or a ctx")
        self.dev = step.dev if step is not None else torch.device("cuda", ctx.device if device is None else device)
        self.ctx = ctx
        h = C.c_void_p()
        handle = (C.c_ubyte * 64)()
        ctx.check(ctx.lib.cat_p2p_create(ctx.handle, self.rank, self.world, C.byref(h), handle))
        self.handle = h
        hdev = self.dev if dist.get_backend(group) == "nccl" else torch.device("cpu")   # gloo: host tensors
        mine = torch.tensor(list(bytes(handle)), dtype=torch.uint8, device=hdev)
        parts = [torch.empty(64, dtype=torch.uint8, device=hdev) for _ in range(self.world)]
        dist.all_gather(parts, mine, group=group)
        buf = (C.c_ubyte * (64 * self.world)).from_buffer_copy(b"".join(bytes(t.cpu().numpy().tobytes()) for t in parts))
        rc = ctx.lib.cat_p2p_connect(self.handle, buf)
        dist.barrier(group)                                 # every rank has mapped every buffer before step 1
        ctx.check(rc)
        self._mode = _P2P_LOCKSTEP

    def _set_mode(self, mode):
        if mode != self._mode:
            self.ctx.check(self.ctx.lib.cat_p2p_set_mode(self.handle, mode))
            self._mode = mode

    def launch(self, winner=None, to_host=False, pipelined=False):
        """Enqueue one step on torch's current stream; ``winner`` (int64[2] device) receives [cost bits, idx];
        ``to_host`` writes it into the step's device-mapped host buffer instead (``step.result_host()``).
        ``pipelined``: the buffer receives the job-wide winner of the PREVIOUS step (see the class docstring)."""
        import ctypes as C
        st = self.step
        wp = st.host_winner().ctypes.data if to_host else (st.winner if winner is None else winner).data_ptr()
        self._set_mode(_P2P_PIPELINED if pipelined else _P2P_LOCKSTEP)
        self.ctx.set_stream(self.torch.cuda.current_stream(st.dev).cuda_stream)
        self.ctx.lib.cat_set_workspace(self.ctx.handle, st._ws)
        fn = self.ctx.lib.cat_mpc_step_p2p_dev_f64 if st.dtype == np.float64 else self.ctx.lib.cat_mpc_step_p2p_dev_f32
        self.ctx.check(fn(self.ctx.handle, self.handle, C.byref(st.params), st.problems.data_ptr(),
                                None if st.controls is None else st.controls.data_ptr(), st.p_ref.data_ptr(),
                                wp, wp + 8, None if st.costs is None else st.costs.data_ptr()))

    def flush(self, winner=None):
        """Enqueue the fold of the most recent step's winners (closes a run of pipelined steps)."""
        st = self.step
        wp = (st.winner if winner is None else winner).data_ptr()
        self.ctx.set_stream(self.torch.cuda.current_stream(st.dev).cuda_stream)
        self.ctx.check(self.ctx.lib.cat_p2p_flush_dev(self.ctx.handle, self.handle, wp, wp + 8))

    def check_winner(self, idx):
        """The in-kernel exchange reports a missing peer through the value itself: (NaN, -1)."""
        if idx < 0:
            self.timed_out()                                # synchronises and clears the device flag
            raise RuntimeError("fused winner exchange: a peer's winner did not arrive (rank stalled or gone); "
                               "the ranks are no longer in step")

    def result_host(self):
        i, c = self.step.result_host()
        self.check_winner(int(i[0]))
        return int(i[0]), float(c[0])

    def result(self, winner=None):
        """(best_idx, best_cost) of the whole job (blocks on the stream)."""
        w = (self.step.winner if winner is None else winner).cpu().numpy()
        self.check_winner(int(w[1]))
        return int(w[1]), float(w[:1].view(np.float64)[0])

    def timed_out(self):
        """True if any step since the last call gave up waiting for a peer (clears the flag)."""
        import ctypes as C
        v = C.c_int()
        self.ctx.check(self.ctx.lib.cat_p2p_status(self.handle, C.byref(v)))
        return bool(v.value)

    def close(self):
        if getattr(self, "handle", None) and self.ctx.handle:
            self.ctx.lib.cat_p2p_destroy(self.handle)
        self.handle = None
