#!/usr/bin/env python
"""bench.py -- rollout-steps/s and p50 MPC-step latency of the sampling-MPC hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Headline workload (config.workload = "config3"): 65,536 candidate control sequences x horizon 50 per GPU, explicit
fp64 controls resident in HBM; one "step" = one sampling-MPC solve (rollout of every candidate + argmin).  N > 1:
weak scaling, each rank owns a contiguous shard of an N x 65,536-candidate job and the ranks exchange their 16-byte
winners inside the rollout kernel over NVLink peer memory every step.  One rollout-step = one candidate x one
horizon step.

The same JSON line carries, as extra keys: ``e2e`` (host arrays in, winner out, through the public API),
``configs`` (BASELINE configs 2, 4 and 5, each with its own timing and roofline), ``sensitivity`` (the headline
step off its data-dependent fast path: dt = 0.05, 4x controls, a near-taut pose, with the fraction of warp-steps that
left the warm catenary solve / the fast sine), ``parity`` (an oracle-checked 65,536 x 50 step and the near-tie
margin of the argmin), ``roofline`` (executed work against measured peaks, the SURVEY's 972-flop convention beside
it), ``peaks``, ``cpu_baseline``.  Single-GPU legs run at N = 1 only.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

K_PER_GPU = 65536
H = 50
DT_SURVEY = 0.05                      # SURVEY.md 8(d) config 3 (main_fun.py:127); see BASELINE.md on dt = 0.01
FLOPS_PER_ROLLOUT_STEP = 972.0        # DFMA-equivalent fp64 flops, SURVEY.md 8(d) convention
BYTES_PER_ROLLOUT_STEP = 24.0         # 3 fp64 controls read per rollout-step (explicit controls)
BYTES_PER_SHAPE_CONFIG = 2481.0       # 72 B in + 2,400 B final curve + 9 B (C, valid)
METRIC = "rk4_tether_rollout_steps_per_sec"
UNIT = "rollout-steps/s"

_REAL_STDOUT = None


def emit(line):
    """Write the result line to the process's real stdout (see main)."""
    fh = _REAL_STDOUT or sys.stdout
    fh.write(line + "\n")
    fh.flush()


def log(msg):
    sys.stderr.write(f"[bench] {msg}\n")
    sys.stderr.flush()


def config3_inputs(K, H, seed, sigma):
    """Controls ~ N(0, diag(sigma^2)) clipped to 3 sigma, SoA [H][3][K]."""
    rng = np.random.default_rng(seed)
    return np.clip(rng.standard_normal((H, 3, K)), -3, 3) * np.asarray(sigma)[None, :, None]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True).start()
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------
# reference arm: the CPU implementation of the path, all host cores
# ---------------------------------------------------------------------------------------------------------------
def oracle_chunk(args):
    """One candidate chunk of the CPU arm (the oracle port; bench.py may execute oracle/ only here)."""
    from oracle import catenary_oracle as orc
    U, pref = args
    J = orc.mpc_rollout_costs(orc.default_mpc_params(), U, pref)
    i, c = orc.argmin_mpc(J)
    return c, i


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path = the NumPy/SciPy oracle port (the
    reference itself is Python research scripts that cannot travel to this box and contain no MPC),
    on all host cores, each step a bounded sample of config 3 (the full 65,536 candidates on a box with >= 32
    cores, 16,384 otherwise)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    from oracle import catenary_oracle as orc
    sigma = orc.load_model()["scaler_scale"][3:6]
    cores = os.cpu_count() or 1
    K = K_PER_GPU if cores >= 32 else 16384
    U = np.ascontiguousarray(config3_inputs(K, H, 1234, sigma).transpose(2, 0, 1))     # (K,H,3)
    pref = orc.circle_reference(H, orc.default_mpc_params()["dt"])
    chunks = [(U[i::cores], pref) for i in range(cores)]
    orc.equation("dtheta_dt")
    with mp.get_context("fork").Pool(cores) as pool:
        for _ in range(args.warmup):
            pool.map(oracle_chunk, chunks)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(oracle_chunk, chunks)
        dt = time.perf_counter() - t0
    v = K * H * args.steps / dt
    sample = f"{K} of 65536 candidates x H={H} per step (oracle port, NumPy fp64, fork pool over candidate chunks)"
    emit(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                      "data": "synthetic", "config": {"workload": "config3", "candidates": K, "horizon": H,
                                                      "dt": orc.default_mpc_params()["dt"], "sample": sample},
                      "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                      "gpu_launches": 0}))


# ---------------------------------------------------------------------------------------------------------------
# helpers of the GPU arm
# ---------------------------------------------------------------------------------------------------------------
def _events(torch):
    return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def _median_launch_ms(torch, launch, n=20, before=None):
    """Median CUDA-event duration of single launches on torch's current stream."""
    ts = []
    for i in range(n):
        if before:
            before(i)
        a, b = _events(torch)
        a.record()
        launch()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b))
    return statistics.median(ts)


def _loop_ms(torch, launch, n, before=None):
    """ms per launch over n back-to-back launches between two events."""
    torch.cuda.synchronize()
    a, b = _events(torch)
    a.record()
    for i in range(n):
        if before:
            before(i)
        launch()
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / n


def _load_json(*path):
    try:
        with open(os.path.join(ROOT, *path)) as fh:
            return json.load(fh)
    except (OSError, ValueError):
        return {}


def run_ours(args):
    import torch
    import catenary_b200 as cb
    devmod = importlib.import_module(cb.__name__ + ".device")
    dist_mod = importlib.import_module(cb.__name__ + ".distributed")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sigma = np.asarray(cb.dynamics.load_model()["scaler_scale"][3:6])
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    ctx = cb.Context(local)
    K = K_PER_GPU
    peaks_file = _load_json("MEASURED_PEAKS.json")
    hbm_peak = peaks_file.get("hbm_gbs", 6650.0)
    hbm_src = "MEASURED_PEAKS.json" if peaks_file else "fallback (B200_PROFILING.md)"

    def allmax(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # three rotating control sets (3 x 78.6 MB = 236 MB > 126 MB L2): every timed step reads cold inputs
    n_sets = 3
    host_sets = [ctx.pinned_empty((H, 3, K)) for _ in range(n_sets)]
    for s, buf in enumerate(host_sets):
        buf[...] = config3_inputs(K, H, 1234 + 17 * s + 1000 * rank, sigma)
    dev_sets = [torch.from_numpy(b).to(dev) for b in host_sets]
    step = devmod.MpcDeviceStep(p, pref, K, controls=host_sets[0], k_offset=rank * K, device=local, ctx=ctx)

    # N > 1: the winners are exchanged INSIDE the rollout kernel through NVLink peer memory (CUDA IPC slot
    # buffers, distributed.P2PWinnerExchange): one kernel per step, no NCCL on the data path.  The throughput loop
    # runs it pipelined (each step folds the winners of the previous one: no cross-rank wait), the latency loop
    # lock-step (the kernel ends with the job-wide winner of its own step).  If the IPC mapping is refused in this
    # environment the NCCL path (side-stream all-gather + device fold, double-buffered) runs instead;
    # config.exchange says which.
    fused = pipe = None
    if world > 1:
        try:
            fused = dist_mod.P2PWinnerExchange(step)
        except Exception as exc:                                  # noqa: BLE001 -- any IPC failure -> NCCL path
            log(f"fused P2P exchange unavailable ({exc}); using NCCL")
        ok = torch.tensor([1 if fused is not None else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:
            fused = None
            pipe = dist_mod.PipelinedExchange(step)

    def one(i, wait=False):
        """One MPC step; ``wait`` blocks for the job-wide winner of THIS step (the latency loop)."""
        step.set_controls(dev_sets[i % n_sets])
        if fused is not None:
            fused.launch(to_host=wait, pipelined=not wait)
            return fused.result_host() if wait else None
        if pipe is None:
            step.launch(to_host=wait)                       # latency loop: the kernel writes the 16-byte winner
            return step.result_host() if wait else None     # into device-mapped pinned memory (no D2H copy)
        slot = pipe.submit()
        return pipe.result(slot) if wait else None

    def barrier():
        if pipe is not None:
            pipe.drain()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    sampler.start()
    for i in range(max(args.warmup, 3)):
        one(i)
    barrier()
    t_wait = time.time()
    while not sampler.lines and time.time() - t_wait < 5.0:      # nvidia-smi takes ~1 s to print its first sample
        step.launch()                                            # rank-local keep-busy: NOT a collective step (the
        torch.cuda.synchronize()                                 # ranks leave this loop after different trip counts)
    barrier()
    # ---- device-timed throughput: K steps between two events on the launching stream ----
    l0 = ctx.launch_count
    e0, e1 = _events(torch)
    barrier()
    e0.record()
    for i in range(args.steps):
        one(i)
    if fused is not None:
        fused.flush()                                           # the last step's winners fold inside the timed region
    if pipe is not None:
        torch.cuda.current_stream().wait_stream(pipe.side)      # the last exchanges are part of the timed region
    e1.record()
    barrier()
    launches = ctx.launch_count - l0
    if pipe is not None:
        launches += args.steps                                  # k_reduce_winners, one per step, on the side stream
    ms_total = allmax(e0.elapsed_time(e1))
    # the same loop with programmatic dependent launch switched off: every step waits for the previous one to retire
    ctx.set_overlap(False)
    for i in range(3):
        one(i)
    barrier()
    g0, g1 = _events(torch)
    g0.record()
    for i in range(args.steps):
        one(i)
    if fused is not None:
        fused.flush()
    if pipe is not None:
        torch.cuda.current_stream().wait_stream(pipe.side)
    g1.record()
    barrier()
    ms_serial = allmax(g0.elapsed_time(g1)) / args.steps
    ctx.set_overlap(True)
    exchange_check = None
    if world > 1:
        # the in-kernel exchange (lock-step, and pipelined + flush) must agree with the plain NCCL all-gather + host fold
        step.set_controls(dev_sets[0])
        step.launch()
        want = dist_mod.exchange_winner(step.winner)
        if fused is not None:
            fused.launch()
            got = fused.result()
            fused.launch(pipelined=True)
            fused.flush()
            got2 = fused.result()
            assert not fused.timed_out()
            assert got == want and got2 == want, (got, got2, want)
        else:
            got = pipe.result(pipe.submit())
            assert got == want, (got, want)
        exchange_check = "fused lock-step, fused pipelined + flush and NCCL all-gather + host fold give the same winner"
    # ---- for comparison at N > 1: fused lock-step throughput, and the NCCL path (all-gather + device fold on a side
    #      stream, double-buffered; the winner arrives one step late) ----
    nccl_cmp = lock_cmp = None
    if fused is not None:
        barrier()
        f0, f1 = _events(torch)
        f0.record()
        for i in range(args.steps):
            step.set_controls(dev_sets[i % n_sets])
            fused.launch()
        f1.record()
        barrier()
        lk = allmax(f0.elapsed_time(f1)) / args.steps
        lock_cmp = {"ms_per_step": lk, "value": world * K * H / (lk * 1e-3),
                    "what": "same steps with the fused exchange in lock-step mode (round-1 headline path)"}
        pipe2 = dist_mod.PipelinedExchange(step)
        for i in range(5):
            step.set_controls(dev_sets[i % n_sets])
            pipe2.submit()
        pipe2.drain()
        barrier()
        f0, f1 = _events(torch)
        f0.record()
        for i in range(args.steps):
            step.set_controls(dev_sets[i % n_sets])
            pipe2.submit()
        torch.cuda.current_stream().wait_stream(pipe2.side)
        f1.record()
        pipe2.drain()
        barrier()
        nccl_ms = allmax(f0.elapsed_time(f1)) / args.steps
        lat2 = []
        for i in range(100):
            step.set_controls(dev_sets[i % n_sets])
            a, b = _events(torch)
            a.record()
            pipe2.result(pipe2.submit())
            b.record()
            b.synchronize()
            lat2.append(a.elapsed_time(b))
        lat2.sort()
        nccl_cmp = {"ms_per_step": nccl_ms, "value": world * K * H / (nccl_ms * 1e-3), "p50_ms": lat2[50],
                    "what": "same steps through NCCL all_gather + device fold on a side stream (pipelined)"}
    # ---- per-step latency distribution (p50 MPC solve latency, delivery of the 16-byte winner to the host included) ----
    lat = []
    n_lat = min(max(args.steps, 200), 2000)
    barrier()
    for i in range(n_lat):
        a, b = _events(torch)
        a.record()
        r = one(i, wait=True)
        assert r is not None
        b.record()
        b.synchronize()
        lat.append(a.elapsed_time(b))
    clocks = sampler.stop()
    ms_per_step = ms_total / args.steps
    value = world * K * H / (ms_per_step * 1e-3)
    log(f"value {value:.4e} rollout-steps/s, {ms_per_step:.4f} ms/step, p50 {sorted(lat)[len(lat) // 2]:.4f} ms")

    # ---- end to end through the public host API: host arrays -> H2D -> kernel -> winner on the host ----
    # every rank runs its own shard; N > 1: the winners are exchanged inside the step's kernel (cat_mpc_step_p2p_*),
    # so the call returns the job-wide winner without a host-side collective
    e2e_steps = max(3, min(args.steps, 20))
    ctx.set_stream(None)
    host32 = [ctx.pinned_empty((H, 3, K), np.float32) for _ in range(n_sets)]
    for s in range(n_sets):
        host32[s][...] = host_sets[s]
    fused_host = None
    if fused is not None:
        fused_host = dist_mod.P2PWinnerExchange(ctx=ctx, device=local)

    def e2e_call(U, i, **kw):
        return cb.mpc_sample_step(p, U, pref, layout="H3K", k_offset=rank * K, return_costs=False, ctx=ctx,
                                  p2p=fused_host, **kw)[:2]

    def e2e_rng(i):
        return cb.mpc_sample_step(p, None, pref, K=K, sigma=sigma, seed=i, k_offset=rank * K, return_costs=False,
                                  ctx=ctx, p2p=fused_host)[:2]

    def timed(fn, n=e2e_steps, warm=2):
        for i in range(warm):
            fn(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(n):
            fn(i)
        barrier()
        return allmax((time.perf_counter() - t0) / n)

    f32_s = timed(lambda i: e2e_call(host32[i % n_sets], i))
    f64_s = timed(lambda i: e2e_call(host_sets[i % n_sets], i))
    rng_s = timed(e2e_rng)
    small = 128 + pref.nbytes
    e2e = {"value": world * K * H / f32_s, "unit": UNIT, "h2d_bytes_per_step": host32[0].nbytes + small,
           "d2h_bytes_per_step": 16, "ms_per_step": f32_s * 1e3,
           "h2d_gbs_per_rank": host32[0].nbytes / f32_s / 1e9, "h2d_gbs_aggregate": world * host32[0].nbytes / f32_s / 1e9,
           "note": "mpc_sample_step (cat_mpc_step_f64): explicit candidates from pinned host memory, stored as "
                   "float32 (CAT_MPC_CONTROLS_F32), fp64 arithmetic, winner back on the host"
                   + ("; job-wide winner exchanged inside the kernel (cat_mpc_step_p2p_f64)" if fused_host else ""),
           "f64_controls": {"value": world * K * H / f64_s, "ms_per_step": f64_s * 1e3,
                            "h2d_bytes_per_step": host_sets[0].nbytes + small,
                            "h2d_gbs_aggregate": world * host_sets[0].nbytes / f64_s / 1e9},
           "device_sampled_controls": {"value": world * K * H / rng_s, "ms_per_step": rng_s * 1e3,
                                       "h2d_bytes_per_step": small}}
    log(f"e2e f32-wire {f32_s * 1e3:.3f} ms, f64 {f64_s * 1e3:.3f} ms, device-sampled {rng_s * 1e3:.3f} ms")

    # ---- BASELINE config 4 (4,194,304 x H = 100 sharded over the ranks) and config 5 (4,096 x 4,096 x 50 fleet) ----
    configs = {}
    K4, H4 = 4194304, 100
    lo4, hi4 = dist_mod.shard_range(K4, rank, world)
    pref4 = cb.circle_reference(H4, p["dt"])
    st4 = devmod.MpcDeviceStep(p, pref4, hi4 - lo4, sigma=sigma, seed=99, k_offset=lo4, device=local, ctx=ctx)
    n4 = 5
    c4 = {"K_total": K4, "H": H4, "controls": "device-sampled (Philox + inverse-CDF table), keyed by global index"}
    if world == 1:
        for _ in range(2):
            st4.launch()
        ms4 = _loop_ms(torch, st4.launch, n4)
        c4.update(ms_per_step=ms4, exchange="none")
        w4 = st4.result()
        c4["best_idx"] = int(w4[0][0])
    else:
        fx4 = dist_mod.P2PWinnerExchange(st4) if fused is not None else None
        if fx4 is not None:
            for _ in range(2):
                fx4.launch()
            barrier()
            a, b = _events(torch)
            a.record()
            for _ in range(n4):
                fx4.launch(to_host=True)
                win = fx4.result_host()                          # the winner is consumed every step
            b.record()
            barrier()
            c4["fused_ms_per_step"] = allmax(a.elapsed_time(b)) / n4
            c4["best_idx"] = win[0]
            barrier()
            a, b = _events(torch)
            a.record()
            for _ in range(n4):
                fx4.launch()                                     # winners stay on the device: no host round trip
            b.record()
            barrier()
            c4["fused_back_to_back_ms_per_step"] = allmax(a.elapsed_time(b)) / n4
        for _ in range(2):
            st4.launch()
            want4 = dist_mod.exchange_winner(st4.winner)
        barrier()
        a, b = _events(torch)
        a.record()
        for _ in range(n4):
            st4.launch()
            want4 = dist_mod.exchange_winner(st4.winner)
        b.record()
        barrier()
        c4["nccl_ms_per_step"] = allmax(a.elapsed_time(b)) / n4
        if fx4 is not None:
            assert win == want4, (win, want4)
        c4["ms_per_step"] = c4.get("fused_ms_per_step", c4["nccl_ms_per_step"])
        c4["exchange"] = "fused (lock-step, winner consumed every step)" if fx4 is not None else "nccl"
    c4["rollout_steps_per_s"] = K4 * H4 / (c4["ms_per_step"] * 1e-3)
    configs["config4"] = c4
    del st4
    n_prob, K5, H5 = 4096, 4096, 50
    lo5, hi5 = dist_mod.fleet_shard(n_prob, rank, world)
    rng5 = np.random.default_rng(10_000)
    d5 = rng5.normal(size=(n_prob, 3))
    d5 /= np.linalg.norm(d5, axis=1, keepdims=True)
    d5[:, 2] = -np.abs(d5[:, 2])
    poses = d5 * rng5.uniform(0.8, 2.5, n_prob)[:, None]
    probs5 = cb.make_problems(dict(p, P1=poses[lo5:hi5]), hi5 - lo5)
    t5 = (np.arange(H5) + 1) * p["dt"]
    w5 = 2 * np.pi / 10.0
    prefs5 = poses[lo5:hi5, None, :] + np.stack([0.1 * np.cos(w5 * t5) - 0.1, 0.1 * np.sin(w5 * t5),
                                                 np.zeros(H5)], 1)[None]
    st5 = devmod.MpcDeviceStep(p, prefs5, K5, sigma=sigma, seed=7, problems=probs5, problem_offset=lo5,
                               device=local, ctx=ctx)
    for _ in range(2):
        st5.launch()
    barrier()
    a, b = _events(torch)
    a.record()
    for _ in range(3):
        st5.launch()
    b.record()
    barrier()
    ms5 = allmax(a.elapsed_time(b)) / 3
    configs["config5"] = {"problems": n_prob, "K": K5, "H": H5, "ms_per_sweep": ms5, "communication": "none",
                          "rollout_steps_per_s": n_prob * K5 * H5 / (ms5 * 1e-3)}
    del st5
    log(f"config4 {c4['ms_per_step']:.3f} ms/step, config5 {ms5:.3f} ms/sweep")

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    lat.sort()
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": "config3", "candidates_per_gpu": K, "horizon": H, "dt": p["dt"],
                      "dt_survey": DT_SURVEY,
                      "dt_note": "SURVEY 8(d) quotes config 3 at dt = 0.05; the shipped d(gamma)/dt fed back on itself is "
                                 "a difference equation with multiplier 2.9 there (it overflows over 50 steps), so the "
                                 "closed-loop configs run at 0.01 -- sensitivity.dt_0.05 times the step at the SURVEY value",
                      "controls": "explicit fp64 SoA [H][3][K]", "exchange": ("none" if world == 1 else
                                   "fused in the rollout kernel: 16-B winners through NVLink peer memory (CUDA IPC); "
                                   "throughput loop pipelined (each step folds the previous step's winners), latency "
                                   "loop lock-step" if fused is not None else
                                   "nccl all_gather 16 B/rank + device fold on a side stream, double-buffered"),
                      "launch": "one stream, back-to-back launches with programmatic stream serialization: a step's "
                                "candidates start on the SMs its predecessor's early finishers leave idle; each step "
                                "touches its outputs only after the predecessor has completed",
                      "l2": "3 rotating control sets (236 MB) > 126 MB L2"},
           "p50_ms": lat[len(lat) // 2], "p99_ms": lat[int(len(lat) * 0.99)], "latency_samples": len(lat),
           "value_no_overlap": {"value": world * K * H / (ms_serial * 1e-3), "ms_per_step": ms_serial,
                                "what": "the same loop with programmatic dependent launch off (cat_set_overlap(0)): each "
                                        "rollout starts after the previous one has retired -- the round-1 launch mode"},
           "clocks": clocks, "e2e": e2e, "configs": configs, "gpu_launches": int(launches),
           "exchange_check": exchange_check, "fused_lockstep": lock_cmp, "nccl_exchange": nccl_cmp}
    def add_steady(o):
        # the same executed-work accounting over the timed region of `value`: with programmatic dependent launch the
        # launches overlap, so a step costs ms_per_step of the pipe's time, not a lone launch's duration
        r = o["roofline"]
        fpw = r["executed"]["fp64_inst_per_warp_step"]
        rate = (K / 32.0) * H / (ms_per_step * 1e-3) * 32.0 * fpw
        ipw = r["executed"]["inst_per_warp_step"]
        sm_mhz = (clocks or {}).get("sm_mhz") or 0
        r["steady_state"] = {"ms_per_step": ms_per_step, "achieved": rate / 1e12, "frac": rate / (r["peak"] * 1e12),
                             "issue_slot_frac": (K / 32.0) * H / (ms_per_step * 1e-3) * (ipw + fpw)
                                                / (148 * 4 * sm_mhz * 1e6) if sm_mhz else None,
                             "what": "FP64 thread-instructions per second of ONE GPU over the back-to-back loop that "
                                     "`value` times, against the same measured DFMA rate; issue_slot_frac = issue-port "
                                     "cycles (one per instruction, two per FP64 instruction) over the cycles of the "
                                     "592 schedulers -- a MODEL of the issue port, not a counter (a value a little "
                                     "above 1 says the two-cycle charge is slightly high); at ~1.0 the loop is bound by "
                                     "its instruction count, and more resident warps do not help "
                                     "(profiles/r2_ab_minb5.txt)"}
        return o
    if world == 1:
        out.update(add_steady(single_gpu_legs(torch, cb, devmod, ctx, step, p, pref, sigma, host_sets, host32, dev_sets, clocks,
                                   hbm_peak, hbm_src, value, local)))
    else:
        # the roofline of the dominant kernel still belongs on every line; the long single-GPU legs do not
        out.update(add_steady(roofline_leg(torch, ctx, step, dev_sets, clocks, hbm_peak, hbm_src)))
        out["cpu_baseline"] = None
    emit(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def roofline_leg(torch, ctx, step, dev_sets, clocks, hbm_peak, hbm_src):
    """Dominant kernel k_mpc_rollout<double,double,shipped,explicit>: executed work against the measured FP64 rate.
    `frac` is the share of the FP64 pipe the kernel keeps busy over its elapsed time (thread-level FP64 instructions
    per second / measured DFMA instructions per second); the SURVEY's 972-flop accounting is kept beside it as
    `convention` (it charges 6 Newton iterations, k3 and library transcendentals the kernel does not execute)."""
    K, n_sets = K_PER_GPU, len(dev_sets)
    k_ms = _median_launch_ms(torch, step.launch, 20, before=lambda i: step.set_controls(dev_sets[i % n_sets]))
    dfma = ctx.measure_peak("dfma")
    hw = _load_json("profiles", "r2_hw_counters.json").get("k_mpc_rollout_f64_config3", {})
    ipw = hw.get("inst_per_warp_step", 412.0)
    fpw = hw.get("fp64_inst_per_warp_step", 211.0)
    warp_steps = (K / 32.0) * H / (k_ms * 1e-3)
    sm_mhz = (clocks or {}).get("sm_mhz") or 0
    fp64_rate = warp_steps * 32.0 * fpw                     # thread-level FP64 instructions per second
    conv_tf = K * H * FLOPS_PER_ROLLOUT_STEP / (k_ms * 1e-3) / 1e12
    hbm_ach = K * H * BYTES_PER_ROLLOUT_STEP / (k_ms * 1e-3) / 1e9
    traffic = _load_json("profiles", "traffic.json").get("k_mpc_rollout_f64_config3", {}).get("dram_bytes_per_launch")
    roof = {"bound": "fp64_fma", "kernel": "k_mpc_rollout<double,double,shipped,explicit>", "kernel_ms": k_ms,
            "achieved": fp64_rate / 1e12, "peak": dfma / 1e12, "unit": "T FP64 instr/s", "frac": fp64_rate / dfma,
            "traffic": traffic,
            "what": "executed FP64 thread-instructions per second over the measured DFMA issue rate (cat_measure_peak, "
                    "this run): the FP64-pipe share of the launch's elapsed time",
            "executed": {"inst_per_warp_step": ipw, "fp64_inst_per_warp_step": fpw,
                         "issue_slot_frac": warp_steps * (ipw + fpw) / (148 * 4 * sm_mhz * 1e6) if sm_mhz else None,
                         "source": hw.get("source", "SASS count of the round-1 kernel (tools/sass_hist.py); an FP64 "
                                                    "instruction holds the issue port for two cycles "
                                                    "(tools/ubench/fp64_latency.cu)")},
            "hw": {"fp64_pipe_frac_elapsed": hw.get("fp64_pipe_frac_elapsed"), "issue_frac": hw.get("issue_frac"),
                   "cycles_active_over_elapsed": hw.get("cycles_active_over_elapsed"),
                   "source": hw.get("source", "no ncu capture of this library version under profiles/ yet")},
            "convention": {"flops_per_rollout_step": FLOPS_PER_ROLLOUT_STEP, "achieved_tflops": conv_tf,
                           "peak_tflops": 2 * dfma / 1e12, "frac": conv_tf / (2 * dfma / 1e12),
                           "note": "SURVEY.md 8(d) algorithmic accounting (972 DFMA-equivalent flops per rollout-step); "
                                   "a throughput-in-convention figure, not pipe utilisation"},
            "hbm": {"achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_ach / hbm_peak,
                    "peak_source": hbm_src}}
    return {"roofline": roof}


def single_gpu_legs(torch, cb, devmod, ctx, step, p, pref, sigma, host_sets, host32, dev_sets, clocks, hbm_peak,
                    hbm_src, value, local):
    """Everything that is measured on one GPU only (N = 1)."""
    K, n_sets, dev = K_PER_GPU, len(dev_sets), f"cuda:{local}"
    out = roofline_leg(torch, ctx, step, dev_sets, clocks, hbm_peak, hbm_src)
    dfma = out["roofline"]["peak"] * 1e12
    ffma, mufu = ctx.measure_peak("ffma"), ctx.measure_peak("mufu")
    out["peaks"] = {"dfma_per_s": dfma, "ffma_per_s": ffma, "mufu_ex2_per_s": mufu, "fp64_tflops": 2 * dfma / 1e12,
                    "fp32_tflops": 2 * ffma / 1e12, "hbm_gbs": hbm_peak, "hbm_source": hbm_src,
                    "sm_mhz_under_load": (clocks or {}).get("sm_mhz"),
                    "how": "cat_measure_peak: 8 independent chains per thread, 148 x 8 blocks x 256 threads, best of 3"}
    ctx.set_stream(None)

    # ---- throughput with the tail of each launch filled by the next one: two streams, private workspaces ----
    # (65,536 candidates are 3.46 warps per scheduler: a launch costs what its 4-warp schedulers cost and the 3-warp
    # ones idle for the last ~20%; independent back-to-back steps -- fleet sweeps, MPPI iterations, this loop -- can
    # overlap that tail.  `value` and `p50_ms` stay single-stream.)
    sts = [devmod.MpcDeviceStep(p, pref, K, controls=host_sets[0][:, :, :256], device=local, ctx=ctx,
                                private_workspace=True) for _ in range(2)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    n_ov = 200

    def ov_loop(n):
        for i in range(n):
            j = i & 1
            sts[j].set_controls(dev_sets[i % n_sets])
            with torch.cuda.stream(streams[j]):
                sts[j].launch()
    ov_loop(10)
    torch.cuda.synchronize()
    a, b = _events(torch)
    a.record()
    for s in streams:
        s.wait_stream(torch.cuda.current_stream())
    ov_loop(n_ov)
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    b.record()
    b.synchronize()
    ov_ms = a.elapsed_time(b) / n_ov
    # identical winners: step object j ended on step n_ov - 2 + j; replay those two serially on the single stream
    wins = [int(st.result()[0][0]) for st in sts]
    serial = []
    for j in range(2):
        step.set_controls(dev_sets[(n_ov - 2 + j) % n_sets])
        step.launch()
        serial.append(int(step.result()[0][0]))
    out["value_overlapped"] = {"value": K * H / (ov_ms * 1e-3), "ms_per_step": ov_ms,
                               "what": "independent steps alternating over two CUDA streams (one context, one private "
                                       "workspace per step object): the next launch fills the idle schedulers of the "
                                       "previous launch's tail", "winners_identical_to_serial": wins == serial,
                               "speedup_vs_value": (K * H / (ov_ms * 1e-3)) / value}
    for st in sts:
        st.close()
    ctx.set_stream(None)

    # ---- e2e from ordinary (pageable) NumPy arrays: as they come, and page-locked in place on first sight ----
    pag = [np.array(h) for h in host32]

    def e2e_pag(i):
        cb.mpc_sample_step(p, pag[i % n_sets], pref, layout="H3K", return_costs=False, ctx=ctx)

    def timed(fn, n=10, warm=2):
        for i in range(warm):
            fn(i)
        t0 = time.perf_counter()
        for i in range(n):
            fn(i)
        return (time.perf_counter() - t0) / n
    keep = ctx.REGISTER_MIN_BYTES
    ctx.REGISTER_MIN_BYTES = 1 << 62                        # never register: the driver's pageable path
    pag_s = timed(e2e_pag)
    ctx.REGISTER_MIN_BYTES = keep
    t0 = time.perf_counter()
    for i in range(n_sets):
        e2e_pag(i)                                          # first sight: cudaHostRegister of each array
    reg_first = (time.perf_counter() - t0) / n_sets
    reg_s = timed(e2e_pag, warm=0)
    out["e2e_pageable"] = {"unregistered_ms_per_step": pag_s * 1e3, "registered_ms_per_step": reg_s * 1e3,
                           "first_call_ms": reg_first * 1e3,
                           "what": "the same float32 candidates as ordinary np.ndarray: the driver's pageable staging "
                                   "path vs page-locked in place on first sight (Context.maybe_register, released "
                                   "with the array)"}
    del pag

    # ---- the sampled-controls kernel (what configs 4/5 and the controller run) ----
    st_rng = devmod.MpcDeviceStep(p, pref, K, sigma=sigma, seed=1, device=local, ctx=ctx)
    rng_ms = _median_launch_ms(torch, st_rng.launch, 20)
    rng_loop = _loop_ms(torch, st_rng.launch, 100)
    hwr = _load_json("profiles", "r2_hw_counters.json").get("k_mpc_rollout_f64_rng_config3", {})
    fpw_r = hwr.get("fp64_inst_per_warp_step")
    out["sampled_controls_kernel"] = {
        "kernel": "k_mpc_rollout<double,double,shipped,sampled>", "kernel_ms": rng_ms, "ms_per_step_back_to_back": rng_loop,
        "rollout_steps_per_s": K * H / (rng_loop * 1e-3),
        "roofline": {"bound": "fp64_fma", "unit": "T FP64 instr/s", "peak": dfma / 1e12,
                     "achieved": (K * H * fpw_r / (rng_ms * 1e-3) / 1e12) if fpw_r else None,
                     "frac": (K * H * fpw_r / (rng_ms * 1e-3) / dfma) if fpw_r else None,
                     "inst_per_warp_step": hwr.get("inst_per_warp_step"), "fp64_inst_per_warp_step": fpw_r,
                     "hw": {"fp64_pipe_frac_elapsed": hwr.get("fp64_pipe_frac_elapsed"), "issue_frac": hwr.get("issue_frac")},
                     "source": hwr.get("source", "no ncu capture of this library version under profiles/ yet"),
                     "convention_frac": K * H * FLOPS_PER_ROLLOUT_STEP / (rng_ms * 1e-3) / (2 * dfma)}}
    del st_rng

    # ---- fp32 fast mode (north_star: 1e-4 relative): same workload, float controls, FFMA + MUFU roofline ----
    U32 = torch.from_numpy(host32[0]).to(dev)
    step32 = devmod.MpcDeviceStep(p, pref, K, controls=host_sets[0][:, :, :256], device=local, ctx=ctx, dtype=np.float32)
    step32.set_controls(U32)
    k32_ms = _median_launch_ms(torch, step32.launch, 20)
    sps32 = K * H / (k32_ms * 1e-3)
    # back to back like `value` (programmatic dependent launch), over four rotating control sets (157 MB > L2)
    sets32 = [U32] + [torch.from_numpy(h).to(dev) for h in host32[1:]]
    while len(sets32) < 4:
        sets32.append(U32.clone())
    turn = [0]

    def launch32():
        step32.set_controls(sets32[turn[0] % len(sets32)])
        turn[0] += 1
        step32.launch()
    k32_loop = _loop_ms(torch, launch32, 100)
    step32.set_controls(U32)
    step32.launch()
    step.set_controls(dev_sets[0])
    step.launch()
    out["fp32_fast_mode"] = {
        "kernel": "k_mpc_rollout<float,float,shipped,explicit>", "kernel_ms": k32_ms, "rollout_steps_per_s": sps32,
        "ms_per_step_back_to_back": k32_loop, "rollout_steps_per_s_back_to_back": K * H / (k32_loop * 1e-3),
        "ffma": {"achieved_tflops": sps32 * 232.0 / 1e12, "peak_tflops": 2 * ffma / 1e12, "frac": sps32 * 232.0 / (2 * ffma)},
        "mufu": {"achieved_gops": sps32 * 29.0 / 1e9, "peak_gops": mufu / 1e9, "frac": sps32 * 29.0 / mufu},
        "convention": "232 flops + 29 MUFU per rollout-step (SURVEY.md 8d)",
        "best_idx_matches_f64": bool(int(step32.result()[0][0]) == int(step.result()[0][0]))}
    # fp64 arithmetic on float32-stored candidates (the e2e wire format), device-resident
    stw = devmod.MpcDeviceStep(p, pref, K, controls=host_sets[0][:, :, :256], device=local, ctx=ctx,
                               controls_dtype=np.float32)
    stw.set_controls(U32)
    out["f32_stored_controls_kernel_ms"] = _median_launch_ms(torch, stw.launch, 20)
    del step32, stw, U32, sets32

    # ---- sensitivity: the headline step off its data-dependent fast path ----
    sens = {}
    warp_steps = (K // 32) * H
    cases = {"baseline": (dict(), 1.0),
             "dt_0.05": (dict(dt=0.05), 1.0),
             "controls_x4": (dict(), 4.0),
             "dt_0.05_controls_x4": (dict(dt=0.05), 4.0),
             "near_taut_pose": (dict(P1=(1.9, 1.9, -0.9)), 1.0)}
    for name, (over, scale) in cases.items():
        pp = cb.default_mpc_params(**over)
        sts_ = devmod.MpcDeviceStep(pp, cb.circle_reference(H, pp["dt"], center=pp["P1"]), K,
                                    controls=host_sets[0][:, :, :256], device=local, ctx=ctx)
        sets = dev_sets if scale == 1.0 else [d * scale for d in dev_sets[:2]]
        ms = _median_launch_ms(torch, sts_.launch, 12, before=lambda i: sts_.set_controls(sets[i % len(sets)]))
        ctx.stats(enable=True, reset=True)
        sts_.set_controls(sets[0])
        sts_.launch()
        torch.cuda.synchronize()
        cnt = ctx.stats(enable=False, reset=True)
        sens[name] = {"ms_per_step": ms, "rollout_steps_per_s": K * H / (ms * 1e-3),
                      "solve_cold_frac_of_warp_steps": cnt["solve_cold"] / warp_steps,
                      "sine_cold_entries_per_warp_step": cnt["sine_cold"] / warp_steps,
                      "finite_winner": bool(np.isfinite(sts_.result()[1][0]))}
        del sts_, sets
    sens["what"] = ("config 3 launched with the stated change; cold fractions from cat_mpc_stats (warp-level entries into "
                    "the catenary solve's iteration loop / the full-range sine, per warp and horizon step), counted in a "
                    "separate launch so the atomics do not touch the timing")
    out["sensitivity"] = sens

    # ---- BASELINE config 2: 1,048,576 tether configs x 100 points ----
    n2 = 1 << 20
    A2, B2, L2, th2, ga2 = config2_inputs(n2, 0)
    tA, tB, tth, tga = (torch.as_tensor(x, device=dev) for x in (A2, B2, th2, ga2))
    tL = torch.full((1,), 3.0, dtype=torch.float64, device=dev)
    o2 = torch.empty((n2, 100, 3), dtype=torch.float64, device=dev)
    C2 = torch.empty(n2, dtype=torch.float64, device=dev)
    v2 = torch.empty(n2, dtype=torch.uint8, device=dev)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)

    def shape_launch(which=8):
        ctx.check(ctx.lib.cat_shape_dev_f64(ctx.handle, n2, tA.data_ptr(), tB.data_ptr(), tL.data_ptr(), 0,
                                            tth.data_ptr(), tga.data_ptr(), 100, which, o2.data_ptr(), C2.data_ptr(),
                                            v2.data_ptr()))
    s_ms = _median_launch_ms(torch, shape_launch, 10)
    valid_frac = float((v2 == 3).float().mean())
    # all four transform_catenary outputs (what the reference's call returns): 4 x 2,400 B written per configuration
    del o2
    o2 = torch.empty((4, n2, 100, 3), dtype=torch.float64, device=dev)
    s4_ms = _median_launch_ms(torch, lambda: shape_launch(15), 10)
    ctx.set_stream(None)
    del o2, tA, tB, tth, tga
    out_h = None
    t0 = time.perf_counter()
    (fin,), Ch, vh = cb.catenary_shapes(A2, B2, 3.0, th2, ga2, ctx=ctx)
    first = time.perf_counter() - t0
    t0 = time.perf_counter()
    (fin,), Ch, vh = cb.catenary_shapes(A2, B2, 3.0, th2, ga2, ctx=ctx)
    e2e2 = time.perf_counter() - t0
    pin_out = ctx.pinned_empty((1, n2, 100, 3))
    cb.catenary_shapes(A2, B2, 3.0, th2, ga2, ctx=ctx, out=pin_out)
    t0 = time.perf_counter()
    cb.catenary_shapes(A2, B2, 3.0, th2, ga2, ctx=ctx, out=pin_out)
    e2e2p = time.perf_counter() - t0
    tr2 = _load_json("profiles", "traffic.json").get("k_shape_f64_config2", {}).get("dram_bytes_per_launch")
    ach2 = n2 * BYTES_PER_SHAPE_CONFIG / (s_ms * 1e-3) / 1e9
    out["configs_single_gpu"] = {"config2": {
        "configs": n2, "points": 100, "kernel": "k_shape<double, final>", "kernel_ms": s_ms,
        "configs_per_s": n2 / (s_ms * 1e-3), "valid_frac": valid_frac,
        "inputs": "SURVEY 8(d) config 2: 1% taut, 1% near-vertical, theta, gamma ~ U[-30, 30] deg, L = 3",
        "roofline": {"bound": "hbm", "achieved": ach2, "peak": hbm_peak, "unit": "GB/s", "frac": ach2 / hbm_peak,
                     "traffic": tr2, "bytes_per_config": BYTES_PER_SHAPE_CONFIG, "peak_source": hbm_src},
        "all_four_curves": {"kernel": "k_shape<double, run-time mask, two stores per point>", "kernel_ms": s4_ms,
                            "configs_per_s": n2 / (s4_ms * 1e-3), "bytes_per_config": BYTES_PER_SHAPE_CONFIG + 7200.0,
                            "achieved_gbs": n2 * (BYTES_PER_SHAPE_CONFIG + 7200.0) / (s4_ms * 1e-3) / 1e9,
                            "hbm_frac": n2 * (BYTES_PER_SHAPE_CONFIG + 7200.0) / (s4_ms * 1e-3) / 1e9 / hbm_peak},
        "e2e": {"ms": e2e2 * 1e3, "first_call_ms": first * 1e3, "configs_per_s": n2 / e2e2,
                "h2d_bytes": int(A2.nbytes + B2.nbytes + th2.nbytes + ga2.nbytes + 8),
                "d2h_bytes": int(fin.nbytes + Ch.nbytes + vh.nbytes),
                "d2h_gbs": (fin.nbytes + Ch.nbytes + vh.nbytes) / e2e2 / 1e9,
                "what": "catenary_shapes(A, B, 3.0, theta, gamma) on NumPy arrays: chunked two-lane H2D / kernel / D2H "
                        "pipeline into a freshly allocated (n, 100, 3) result (page faults + pageable D2H)",
                "into_pinned_out": {"ms": e2e2p * 1e3, "configs_per_s": n2 / e2e2p, "d2h_gbs": fin.nbytes / e2e2p / 1e9,
                                    "what": "the same call writing into a reused page-locked buffer (out=ctx.pinned_empty)"}}}}
    del fin, Ch, vh, out_h, pin_out

    # ---- closed loop: the receding-horizon controller (device-resident, one CUDA graph per control step) ----
    closed = {}
    for mode in ("argmin", "mppi"):
        n_ctrl = 60
        rh = cb.RecedingHorizonController(cb.default_mpc_params(w_trk=1e3, w_du=1e-6),
                                          cb.circle_reference(n_ctrl + H, p["dt"], period=4.0), K, H,
                                          sigma=(40.0, 40.0, 20.0), u_max=(300.0, 300.0, 300.0), mode=mode, lam=0.02,
                                          seed=3, max_steps=n_ctrl, ctx=ctx)
        rh.run(10)
        ctx.sync()
        ctx.timer_start()
        rh.run(n_ctrl - 10)
        ms = ctx.timer_stop_ms() / (n_ctrl - 10)
        lg = rh.read()["log"][0]
        rh.close()
        closed[mode] = {"ms_per_control_step": ms, "rollout_steps_per_s": K * H / (ms * 1e-3),
                        "kernels_per_step": 3 if mode == "mppi" else 2, "host_bytes_per_step": 0,
                        "final_tracking_error_m": float(np.linalg.norm(
                            lg[-1, :3] - cb.circle_reference(n_ctrl + H, p["dt"], period=4.0)[n_ctrl - 1]))}
    out["closed_loop"] = closed

    # ---- config 1 (the reference's own CPU-runnable case): 5,000-step trajectory, host arrays in and out ----
    feat = importlib.import_module(cb.__name__ + ".features")
    cols1 = feat.config1_trajectory(5000)
    feat.config1_pipeline(cols1, ctx=ctx)
    t0 = time.perf_counter()
    for _ in range(5):
        feat.config1_pipeline(cols1, ctx=ctx)
    c1_s = (time.perf_counter() - t0) / 5
    config1 = {"steps": 5000, "gpu_ms": c1_s * 1e3, "gpu_steps_per_s": 5000 / c1_s,
               "what": "extract_features + scaler + RK4 of both ODEs + per-step solve_catenary + final curve "
                       "(T,100,3), NumPy in / NumPy out through the C ABI"}

    # ---- parity: an oracle-checked 65,536 x 50 step and the near-tie margin of the argmin ----
    from oracle import catenary_oracle as orc
    from oracle import near_ties
    po = orc.default_mpc_params()
    Uo = np.ascontiguousarray(host_sets[0].transpose(2, 0, 1))               # (K,H,3)
    t0 = time.perf_counter()
    Jo = orc.mpc_rollout_costs(po, Uo, pref)
    t_or = time.perf_counter() - t0
    bi, bc, J = cb.mpc_sample_step(p, host_sets[0], pref, layout="H3K", ctx=ctx)
    srt = np.sort(Jo)
    probes = []
    Kp = 8192
    for case in near_ties.plant_near_ties(po, Uo[:Kp], pref, gaps=(1e-10, 1e-12, 1e-13, 1e-14, 1e-15)):
        gi, _, _ = cb.mpc_sample_step(p, case["U"], pref, return_costs=False, ctx=ctx)
        probes.append((case["gap"], bool(gi == case["want"])))
    mg = near_ties.margin(probes)
    out["parity"] = {"workload": "config 3, 65,536 x 50, explicit fp64 controls, every cost against the NumPy oracle",
                     "max_rel_err_J": float(np.abs(J / Jo - 1).max()), "index_exact": bool(bi == orc.argmin_mpc(Jo)[0]),
                     "best_second_rel_gap": float((srt[1] - srt[0]) / srt[0]),
                     "min_gap_tested": min(g for g, _ in probes), "min_gap_exact": mg["min_gap_exact"],
                     "first_mismatch_gap": mg["first_mismatch_gap"], "near_tie_cases": mg["cases"],
                     "near_ties": "pairs planted by oracle/near_ties.py in an 8,192-candidate batch: oracle costs a "
                                  "stated relative gap apart, the better one at the higher index in another block",
                     "oracle_s": t_or}

    # ---- CPU baseline: the oracle port on this box's host cores, bounded sample ----
    Ks = 16384
    Uc = Uo[:Ks]
    t0 = time.perf_counter()
    reps = 0
    while time.perf_counter() - t0 < 8.0:
        orc.mpc_rollout_costs(po, Uc, pref)
        reps += 1
    cpu_v = reps * Ks * H / (time.perf_counter() - t0)
    # config 1 on the CPU, the reference's way: per-step Python loops (rk4_integration with one predict per stage,
    # brentq per row, transform_catenary's per-point list comprehensions), bounded samples, extrapolated per step
    ns = 400
    Xo = orc.scale_features(orc.extract_features({k: v[:ns] for k, v in cols1.items()}))
    t1 = time.perf_counter()
    th_o = orc.rk4_integration(orc.EquationModel("dtheta_dt"), Xo, cols1["Time"][:ns], cols1["Theta"][0])
    ga_o = orc.rk4_integration(orc.EquationModel("dgamma_dt"), Xo, cols1["Time"][:ns], cols1["Gamma"][0])
    t_rk4 = (time.perf_counter() - t1) / ns
    P1o = np.stack([cols1[f"robot_cable_attach_point {a}"][:ns] for a in "XYZ"], 1) / 1000
    t1 = time.perf_counter()
    orc.solve_catenary_brentq(np.hypot(P1o[:, 0], P1o[:, 1]), P1o[:, 2], 3.0)
    t_solve = (time.perf_counter() - t1) / ns
    t1 = time.perf_counter()
    cat_o = orc.Catenary(3.0)
    for i in range(40):
        orc.transform_catenary(np.zeros(3), P1o[i], cat_o, th_o[i], ga_o[i])
    t_shape = (time.perf_counter() - t1) / 40
    config1["cpu_port_steps_per_s"] = 1.0 / (t_rk4 + t_solve + t_shape)
    config1["cpu_port_us_per_step"] = {"rk4_both_odes": t_rk4 * 1e6, "solve_catenary_brentq": t_solve * 1e6,
                                       "transform_catenary": t_shape * 1e6}
    config1["cpu_sample"] = f"{ns} RK4 steps, {ns} brentq solves, 40 transform_catenary calls, 1 core, extrapolated"
    out["config1"] = config1
    out["cpu_baseline"] = {"value": cpu_v, "unit": UNIT, "cores": 1, "kind": "port", "host_cores": os.cpu_count(),
                           "sample": f"{reps} x ({Ks} of 65536 candidates x H={H}), vectorised NumPy fp64 oracle, 1 thread"}
    return out


def config2_inputs(n, seed):
    """P0 = 0, direction uniform on the sphere, d ~ U[0.3, 2.85], 1% taut (d ~ U[3.0, 3.2]),
    1% near-vertical (l < 1e-6), theta, gamma ~ U[-30, 30] deg, L = 3."""
    rng = np.random.default_rng(seed)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    dist = rng.uniform(0.3, 2.85, n)
    n1 = max(n // 100, 1)
    dist[:n1] = rng.uniform(3.0, 3.2, n1)
    d[n1:2 * n1, :2] *= 1e-7
    d[n1:2 * n1] /= np.linalg.norm(d[n1:2 * n1], axis=1, keepdims=True)
    perm = rng.permutation(n)                                   # the sub-populations spread over the batch
    A = np.zeros((n, 3))
    B = np.ascontiguousarray((d * dist[:, None])[perm])
    return A, B, np.full(n, 3.0), np.radians(rng.uniform(-30, 30, n)), np.radians(rng.uniform(-30, 30, n))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    # stdout carries exactly ONE line (the JSON): libraries that write to fd 1 (NCCL prints its version banner there
    # on communicator creation) are sent to stderr, and the JSON goes to the saved descriptor at the end.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        if args.steps > 20:
            args.steps = 20                     # each CPU step is ~1 s of all-core work; keep the arm to minutes
        args.warmup = min(args.warmup, 2)
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
