"""Multi-GPU sharding of the sampling-MPC step: one process per GPU (torch.distributed), candidates split into
contiguous shards, ONE (cost, index) exchange per step.

NCCL has no MINLOC and an fp64 cost + int64 index do not pack into one 64-bit min, so every rank all-gathers
the 16-byte winners and folds them in rank order with the lowest-global-index tie-break
(cat_reduce_winners_f64) -- identical on every rank and identical to np.argmin over the concatenated costs.
The fleet configuration (independent problems) shards problems and needs no communication.
"""
from __future__ import annotations

import numpy as np

from .mpc import reduce_winners


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
    """Fused multi-GPU step: ONE kernel per step per rank.  The rollout's last block writes this
    rank's (cost, global idx) into every peer's slot buffer over NVLink (CUDA IPC mappings), waits for the peers'
    winners of the same step and folds them in rank order -- when the kernel ends, ``step.winner`` holds the
    JOB-WIDE winner on every rank (same value as ``exchange_winner``), with no NCCL call and no second launch.
    The process group is used once, at construction, to swap the 64-byte IPC handles."""

    def __init__(self, step, group=None):
        import ctypes as C

        import torch
        import torch.distributed as dist
        self.step, self.torch = step, torch
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if step.n != 1:
            raise ValueError("the fused exchange handles one problem per launch")
        ctx = step.ctx
        self.ctx = ctx
        h = C.c_void_p()
        handle = (C.c_ubyte * 64)()
        ctx.check(ctx.lib.cat_p2p_create(ctx.handle, self.rank, self.world, C.byref(h), handle))
        self.handle = h
        hdev = step.dev if dist.get_backend(group) == "nccl" else torch.device("cpu")   # gloo: host tensors
        mine = torch.tensor(list(bytes(handle)), dtype=torch.uint8, device=hdev)
        parts = [torch.empty(64, dtype=torch.uint8, device=hdev) for _ in range(self.world)]
        dist.all_gather(parts, mine, group=group)
        buf = (C.c_ubyte * (64 * self.world)).from_buffer_copy(b"".join(bytes(t.cpu().numpy().tobytes()) for t in parts))
        rc = ctx.lib.cat_p2p_connect(self.handle, buf)
        dist.barrier(group)                                 # every rank has mapped every buffer before step 1
        ctx.check(rc)
        self._fn = ctx.lib.cat_mpc_step_p2p_dev_f64 if step.dtype == np.float64 else ctx.lib.cat_mpc_step_p2p_dev_f32

    def launch(self, winner=None, to_host=False):
        """Enqueue one step on torch's current stream; ``winner`` (int64[2] device) receives [cost bits, idx];
        ``to_host`` writes it into the step's device-mapped host buffer instead (``step.result_host()``)."""
        import ctypes as C
        st = self.step
        wp = st.host_winner().ctypes.data if to_host else (st.winner if winner is None else winner).data_ptr()
        self.ctx.set_stream(self.torch.cuda.current_stream(st.dev).cuda_stream)
        self.ctx.check(self._fn(self.ctx.handle, self.handle, C.byref(st.params), st.problems.data_ptr(),
                                None if st.controls is None else st.controls.data_ptr(), st.p_ref.data_ptr(),
                                wp, wp + 8, None if st.costs is None else st.costs.data_ptr()))

    def result_host(self):
        i, c = self.step.result_host()
        return int(i[0]), float(c[0])

    def result(self, winner=None):
        """(best_idx, best_cost) of the whole job (blocks on the stream)."""
        w = (self.step.winner if winner is None else winner).cpu().numpy()
        return int(w[1]), float(w[:1].view(np.float64)[0])

    def timed_out(self):
        import ctypes as C
        v = C.c_int()
        self.ctx.check(self.ctx.lib.cat_p2p_status(self.handle, C.byref(v)))
        return bool(v.value)

    def close(self):
        if getattr(self, "handle", None) and self.ctx.handle:
            self.ctx.lib.cat_p2p_destroy(self.handle)
        self.handle = None
