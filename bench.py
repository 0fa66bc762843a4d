#!/usr/bin/env python
"""bench.py -- rollout-steps/s and p50 MPC-step latency of the sampling-MPC hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (config.workload = "config3"): 65,536 candidate control sequences x horizon 50 per GPU, explicit fp64
controls; one "step" = one sampling-MPC solve (rollout of every candidate + argmin).  N > 1: weak scaling, each
rank owns a contiguous shard of an N x 65,536-candidate job and the ranks exchange their 16-byte winners over
NCCL every step.  One rollout-step = one candidate x one horizon step.
"""
from __future__ import annotations

import argparse
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
FLOPS_PER_ROLLOUT_STEP = 972.0        # DFMA-equivalent fp64 flops, SURVEY.md 8(d)
BYTES_PER_ROLLOUT_STEP = 24.0         # 3 fp64 controls read per rollout-step (explicit controls)
# Instructions the fp64 kernel actually executes per warp and horizon step on its common path (SASS of
# k_mpc_rollout<double,false,false,false>, counted with tools/sass_hist.py): 400 in all, 211 of
# them on the FP64 pipe.  An FP64 instruction holds the issue port for two cycles and nothing issues beside it
# (tools/ubench/fp64_latency.cu), so a warp-step costs 400 + 211 issue slots.
EXEC_INSTR_PER_WARP_STEP = 400.0
EXEC_FP64_PER_WARP_STEP = 211.0
METRIC = "rk4_tether_rollout_steps_per_sec"
UNIT = "rollout-steps/s"


_REAL_STDOUT = None


def emit(line):
    """Write the result line to the process's real stdout (see main)."""
    fh = _REAL_STDOUT or sys.stdout
    fh.write(line + "\n")
    fh.flush()


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


def oracle_chunk(args):
    """One candidate chunk of the CPU arm (the oracle port; bench.py may execute oracle/ only here)."""
    from oracle import catenary_oracle as orc
    U, pref = args
    J = orc.mpc_rollout_costs(orc.default_mpc_params(), U, pref)
    i = int(np.argmin(J))
    return float(J[i]), i


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path = the NumPy/SciPy oracle port (the
    reference itself is Python research scripts that cannot travel to this box and contain no MPC),
    on all host cores, each step a bounded sample of config 3."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    from oracle import catenary_oracle as orc
    sigma = orc.load_model()["scaler_scale"][3:6]
    cores = os.cpu_count() or 1
    K = 16384                                   # bounded sample: 16,384 of the 65,536 candidates, full horizon
    U = np.ascontiguousarray(config3_inputs(K, H, 1234, sigma).transpose(2, 0, 1))     # (K,H,3)
    pref = orc.circle_reference(H, 0.01)
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
                                                      "sample": sample},
                      "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                      "gpu_launches": 0}))


def _executed(K, H, k_ms, dfma_per_s, clocks):
    """What the dominant kernel executes, next to the algorithmic convention of `roofline`: FP64-pipe share
    (thread-level FP64 instructions per second over the measured DFMA rate) and issue-slot share (slots per
    second over sub-partitions x SM clock sampled under load).  Explains where the time goes; not a claim."""
    warp_steps = (K / 32.0) * H / (k_ms * 1e-3)
    sm_mhz = (clocks or {}).get("sm_mhz") or 0
    slots = warp_steps * (EXEC_INSTR_PER_WARP_STEP + EXEC_FP64_PER_WARP_STEP)
    out = {"instr_per_warp_step": EXEC_INSTR_PER_WARP_STEP, "fp64_instr_per_warp_step": EXEC_FP64_PER_WARP_STEP,
           "fp64_pipe_frac": warp_steps * 32.0 * EXEC_FP64_PER_WARP_STEP / dfma_per_s,
           "issue_slot_frac": slots / (148 * 4 * sm_mhz * 1e6) if sm_mhz else None,
           "source": "SASS common path (tools/sass_hist.py), FP64 = 2 issue slots (tools/ubench/fp64_latency.cu)"}
    return out


def _f64_winner(step, controls):
    step.set_controls(controls)
    step.launch()
    return step.result()[0][0]


def run_ours(args):
    import torch
    import catenary_b200 as cb
    import importlib
    devmod = importlib.import_module(cb.__name__ + ".device")
    dist_mod = importlib.import_module(cb.__name__ + ".distributed")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sigma = cb.dynamics.load_model()["scaler_scale"][3:6]
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    ctx = cb.Context(local)
    K = K_PER_GPU
    # three rotating control sets (3 x 78.6 MB = 236 MB > 126 MB L2): every timed step reads cold inputs
    n_sets = 3
    host_sets = [ctx.pinned_empty((H, 3, K)) for _ in range(n_sets)]
    for s, buf in enumerate(host_sets):
        buf[...] = config3_inputs(K, H, 1234 + 17 * s + 1000 * rank, sigma)
    dev_sets = [torch.from_numpy(b).to(f"cuda:{local}") for b in host_sets]
    step = devmod.MpcDeviceStep(p, pref, K, controls=host_sets[0], k_offset=rank * K, device=local, ctx=ctx)

    # N > 1: the winners are exchanged INSIDE the rollout kernel through NVLink peer memory (CUDA IPC slot
    # buffers, distributed.P2PWinnerExchange): one kernel per step, no NCCL on the data path.  If the IPC mapping
    # is refused in this environment the NCCL path (side-stream all-gather + device fold, double-buffered) runs
    # instead; config.exchange says which.
    fused = pipe = None
    if world > 1:
        try:
            fused = dist_mod.P2PWinnerExchange(step)
        except Exception as exc:                                  # noqa: BLE001 -- any IPC failure -> NCCL path
            sys.stderr.write(f"[bench] fused P2P exchange unavailable ({exc}); using NCCL\n")
        ok = torch.tensor([1 if fused is not None else 0], device=f"cuda:{local}")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:
            fused = None
            pipe = dist_mod.PipelinedExchange(step)

    def one(i, wait=False):
        """One MPC step; ``wait`` blocks for the job-wide winner (the latency loop)."""
        step.set_controls(dev_sets[i % n_sets])
        if fused is not None:
            fused.launch(to_host=wait)
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
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        one(i)
    if pipe is not None:
        torch.cuda.current_stream().wait_stream(pipe.side)      # the last exchanges are part of the timed region
    e1.record()
    barrier()
    launches = ctx.launch_count - l0
    if pipe is not None:
        launches += args.steps                                  # k_reduce_winners, one per step, on the side stream
    if world > 1:
        # the in-kernel / pipelined exchange must agree with the plain NCCL all-gather + host fold
        step.launch()
        want = dist_mod.exchange_winner(step.winner)
        if fused is not None:
            fused.launch()
            got = fused.result()
            assert not fused.timed_out()
        else:
            got = pipe.result(pipe.submit())
        assert got == want, (got, want)
    ms_total = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms_total], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    # ---- for comparison at N > 1: the NCCL path (all-gather + device fold on a side stream, double-buffered so
    #      the exchange of step i overlaps the rollout of step i+1; the winner arrives one step late) ----
    nccl_cmp = None
    if fused is not None:
        pipe2 = dist_mod.PipelinedExchange(step)
        for i in range(5):
            step.set_controls(dev_sets[i % n_sets])
            pipe2.submit()
        pipe2.drain()
        dist.barrier()
        torch.cuda.synchronize()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for i in range(args.steps):
            step.set_controls(dev_sets[i % n_sets])
            pipe2.submit()
        torch.cuda.current_stream().wait_stream(pipe2.side)
        f1.record()
        pipe2.drain()
        dist.barrier()
        torch.cuda.synchronize()
        t = torch.tensor([f0.elapsed_time(f1)], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        nccl_ms = float(t.item()) / args.steps
        lat2 = []
        for i in range(100):
            step.set_controls(dev_sets[i % n_sets])
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            pipe2.result(pipe2.submit())
            b.record()
            b.synchronize()
            lat2.append(a.elapsed_time(b))
        lat2.sort()
        nccl_cmp = {"ms_per_step": nccl_ms, "value": world * K * H / (nccl_ms * 1e-3), "p50_ms": lat2[50],
                    "what": "same steps through NCCL all_gather + device fold on a side stream (pipelined)"}
    # ---- per-step latency distribution (p50 MPC solve latency, D2H of the 16-byte winner included) ----
    lat = []
    n_lat = min(max(args.steps, 200), 2000)
    for i in range(n_lat):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        r = one(i, wait=True)
        assert r is not None
        b.record()
        b.synchronize()
        lat.append(a.elapsed_time(b))
    clocks = sampler.stop()
    ms_per_step = ms_total / args.steps
    value = world * K * H / (ms_per_step * 1e-3)
    # ---- end to end through the public host API: pinned host controls -> H2D -> kernel -> D2H winner ----
    # every rank runs its own shard through cat_mpc_step_f64; N > 1 adds the winner exchange on the host
    e2e_steps = max(3, min(args.steps, 20))
    h2d = host_sets[0].nbytes + 128 + pref.nbytes
    ctx.set_stream(None)

    def e2e_one(i, rng=False):
        if rng:
            bi, bc, _ = cb.mpc_sample_step(p, None, pref, K=K, sigma=sigma, seed=i, k_offset=rank * K,
                                           return_costs=False, ctx=ctx)
        else:
            bi, bc, _ = cb.mpc_sample_step(p, host_sets[i % n_sets], pref, layout="H3K", k_offset=rank * K,
                                           return_costs=False, ctx=ctx)
        if world > 1:
            return dist_mod.exchange_winner(dist_mod.pack_winner(bc, bi).to(f"cuda:{local}"))
        return bi, bc

    def timed(rng):
        for i in range(2):
            e2e_one(i, rng)
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            e2e_one(i, rng)
        barrier()
        dt = (time.perf_counter() - t0) / e2e_steps
        if world > 1:
            t = torch.tensor([dt], device=f"cuda:{local}", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return dt

    e2e_s = timed(False)
    rng_s = timed(True)
    e2e = {"value": world * K * H / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 16,
           "ms_per_step": e2e_s * 1e3, "note": "cat_mpc_step_f64 with pinned host controls on every rank",
           "device_sampled_controls": {"value": world * K * H / rng_s, "ms_per_step": rng_s * 1e3,
                                       "h2d_bytes_per_step": 128 + pref.nbytes}}
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    # ---- kernel-alone duration for the roofline (single launches, cold controls) ----
    kt = []
    for i in range(20):
        step.set_controls(dev_sets[i % n_sets])
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step.launch()
        b.record()
        b.synchronize()
        kt.append(a.elapsed_time(b))
    k_ms = statistics.median(kt)
    dfma = ctx.measure_peak("dfma")
    peak_tf = 2 * dfma / 1e12
    ach_tf = K * H * FLOPS_PER_ROLLOUT_STEP / (k_ms * 1e-3) / 1e12
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            peaks = json.load(fh)
    except OSError:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    hbm_ach = K * H * BYTES_PER_ROLLOUT_STEP / (k_ms * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            traffic = json.load(fh)["k_mpc_rollout_f64_config3"]["dram_bytes_per_launch"]
    except (OSError, KeyError, ValueError):
        pass
    roofline = {"bound": "fp64_fma", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": ach_tf / peak_tf, "traffic": traffic, "kernel": "k_mpc_rollout<double,false>",
                "kernel_ms": k_ms, "peak_source": "cat_measure_peak(DFMA) measured in this run x 2 flops",
                "flops_per_rollout_step": FLOPS_PER_ROLLOUT_STEP,
                "executed": _executed(K, H, k_ms, dfma, clocks),
                "hbm": {"achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_ach / hbm_peak,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}}
    # ---- fp32 fast mode (north_star: 1e-4 relative): same workload, float controls, FFMA + MUFU roofline ----
    U32 = torch.from_numpy(host_sets[0].astype(np.float32)).to(f"cuda:{local}")
    step32 = devmod.MpcDeviceStep(p, pref, K, controls=host_sets[0][:, :, :256], device=local, ctx=ctx,
                                  dtype=np.float32)          # small upload; the full fp32 set is swapped in below
    step32.set_controls(U32)
    kt32 = []
    for i in range(20):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step32.launch()
        b.record()
        b.synchronize()
        kt32.append(a.elapsed_time(b))
    k32_ms = statistics.median(kt32)
    ffma, mufu = ctx.measure_peak("ffma"), ctx.measure_peak("mufu")
    sps32 = K * H / (k32_ms * 1e-3)
    fp32 = {"kernel": "k_mpc_rollout<float,false>", "kernel_ms": k32_ms, "rollout_steps_per_s": sps32,
            "ffma": {"achieved_tflops": sps32 * 232.0 / 1e12, "peak_tflops": 2 * ffma / 1e12,
                     "frac": sps32 * 232.0 / (2 * ffma)},
            "mufu": {"achieved_gops": sps32 * 29.0 / 1e9, "peak_gops": mufu / 1e9, "frac": sps32 * 29.0 / mufu},
            "convention": "232 flops + 29 MUFU per rollout-step (SURVEY.md 8d)",
            "best_idx_matches_f64": bool(int(step32.result()[0][0]) == int(_f64_winner(step, dev_sets[0])))}
    # ---- closed loop: the receding-horizon controller (device-resident, one CUDA graph per control step) ----
    ctx.set_stream(None)
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
        log = rh.read()["log"][0]
        rh.close()
        closed[mode] = {"ms_per_control_step": ms, "rollout_steps_per_s": K * H / (ms * 1e-3),
                        "kernels_per_step": 3 if mode == "mppi" else 2, "host_bytes_per_step": 0,
                        "final_tracking_error_m": float(np.linalg.norm(
                            log[-1, :3] - cb.circle_reference(n_ctrl + H, p["dt"], period=4.0)[n_ctrl - 1]))}
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
    # ---- CPU baseline: the oracle port on this box's host cores, bounded sample ----
    from oracle import catenary_oracle as orc
    Ks = 16384
    Uc = np.ascontiguousarray(host_sets[0][:, :, :Ks].transpose(2, 0, 1))
    orc.mpc_rollout_costs(orc.default_mpc_params(), Uc[:256], pref)          # import / lambdify warm-up
    t0 = time.perf_counter()
    reps = 0
    while time.perf_counter() - t0 < 10.0:
        orc.mpc_rollout_costs(orc.default_mpc_params(), Uc, pref)
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
    cpu = {"value": cpu_v, "unit": UNIT, "cores": 1, "kind": "port", "host_cores": os.cpu_count(),
           "sample": f"{reps} x ({Ks} of 65536 candidates x H={H}), vectorised NumPy fp64 oracle, 1 thread"}
    lat.sort()
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": "config3", "candidates_per_gpu": K, "horizon": H, "dt": p["dt"],
                      "controls": "explicit fp64 SoA [H][3][K]", "exchange": ("none" if world == 1 else
                                   "fused in the rollout kernel: 16-B winners through NVLink peer memory (CUDA IPC), "
                                   "rank-order fold in the last block" if fused is not None else
                                   "nccl all_gather 16 B/rank + device fold on a side stream, double-buffered"),
                      "l2": "3 rotating control sets (236 MB) > 126 MB L2"},
           "p50_ms": lat[len(lat) // 2], "p99_ms": lat[int(len(lat) * 0.99)], "latency_samples": len(lat),
           "clocks": clocks, "e2e": e2e, "closed_loop": closed, "config1": config1, "fp32_fast_mode": fp32, "nccl_exchange": nccl_cmp, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu}
    emit(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


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
