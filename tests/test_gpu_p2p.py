"""Fused peer-memory winner exchange on ONE GPU: two ranks (gloo for the handle swap) share
cuda:0, map each other's slot buffers through CUDA IPC and run the one-kernel step -- lock-step, pipelined (winner one
step late + flush), the two interleaved, and the host-pointer entry with device-sampled and with explicit controls.
The two processes' kernels are time-sliced, so this checks the protocol (sequence numbers, ring slots, rank-order fold),
not the latency."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import catenary_oracle as orc

pytestmark = pytest.mark.gpu
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")

WORKER = r"""
import importlib, json, os, sys
sys.path.insert(0, sys.argv[1])
import numpy as np, torch, torch.distributed as dist
import catenary_b200 as cb
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
torch.cuda.set_device(0)
dmod = importlib.import_module(cb.__name__ + ".distributed")
devmod = importlib.import_module(cb.__name__ + ".device")
sigma = cb.dynamics.load_model()["scaler_scale"][3:6]
p = cb.default_mpc_params()
K_total, H = 6000, 12
pref = cb.circle_reference(H, p["dt"])
lo, hi = dmod.shard_range(K_total, rank, world)
SEEDS = (3, 4, 5, 6, 7, 8, 9)
out = {"rank": rank, "winners": [], "pipelined": [], "host": [], "host_explicit": None}
try:
    steps = [devmod.MpcDeviceStep(p, pref, hi - lo, sigma=sigma, seed=s, k_offset=lo, device=0) for s in SEEDS]
    ex = dmod.P2PWinnerExchange(steps[0])
    for st in steps:                      # lock-step: seven steps, every ring slot reused, a fresh job each time
        ex.step = st
        ex.launch()
        out["winners"].append(list(ex.result()))
    # pipelined: step s returns the job-wide winner of step s-1; flush() closes the run
    wbuf = [torch.zeros(2, dtype=torch.int64, device="cuda:0") for _ in SEEDS]
    for j, st in enumerate(steps):
        ex.step = st
        ex.launch(winner=wbuf[j], pipelined=True)
    last = torch.zeros(2, dtype=torch.int64, device="cuda:0")
    ex.flush(winner=last)
    torch.cuda.synchronize()
    out["pipelined_first"] = list(ex.result(winner=wbuf[0]))   # = the winner of the step before the run
    for j in range(1, len(SEEDS)):
        out["pipelined"].append(list(ex.result(winner=wbuf[j])))
    out["pipelined"].append(list(ex.result(winner=last)))
    # modes interleaved, then the host-pointer entry (NumPy in, job-wide winner out)
    ex.step = steps[0]
    ex.launch()
    out["after_mix"] = list(ex.result())
    for s in SEEDS[:3]:
        bi, bc, _ = cb.mpc_sample_step(p, None, pref, K=hi - lo, sigma=sigma, seed=s, k_offset=lo,
                                       return_costs=False, ctx=steps[0].ctx, p2p=ex)
        out["host"].append([bi, bc])
    # explicit controls through the chunked host pipeline + the one-block exchange kernel
    Kx = 80_000
    xlo, xhi = dmod.shard_range(Kx, rank, world)
    Ux = (np.clip(np.random.default_rng(99).normal(size=(Kx, H, 3)), -3, 3) * np.asarray(sigma))[xlo:xhi]
    bi, bc, _ = cb.mpc_sample_step(p, Ux.astype(np.float32), pref, k_offset=xlo, return_costs=False,
                                   ctx=steps[0].ctx, p2p=ex)
    out["host_explicit"] = [bi, bc]
    out["timed_out"] = ex.timed_out()
except Exception as e:                    # IPC refused in this environment, ...
    import traceback
    out["error"] = repr(e) + traceback.format_exc()[-600:]
print("RESULT " + json.dumps(out), flush=True)
dist.barrier()
"""


def test_fused_exchange_two_ranks_one_gpu(cb, tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
        procs.append(subprocess.Popen([sys.executable, str(script), ROOT], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = []
    for pr in procs:
        try:
            so, se = pr.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            pytest.fail("fused exchange workers timed out")
        line = [ln for ln in so.splitlines() if ln.startswith("RESULT ")]
        assert line, se[-2000:]
        outs.append(json.loads(line[-1][7:]))
    if any("error" in o for o in outs):
        pytest.skip(f"CUDA IPC between two processes unavailable here: {[o.get('error') for o in outs]}")
    assert outs[0]["winners"] == outs[1]["winners"] and not outs[0]["timed_out"] and not outs[1]["timed_out"]
    sigma = np.array(orc.load_model()["scaler_scale"][3:6])
    p = orc.default_mpc_params()
    pref = orc.circle_reference(12, p["dt"])
    seeds = (3, 4, 5, 6, 7, 8, 9)
    for seed, (bi, bc) in zip(seeds, outs[0]["winners"]):
        J = orc.mpc_rollout_costs(p, orc.sample_controls(seed, np.arange(6000), 12, sigma), pref)
        assert bi == int(np.argmin(J)) and abs(bc - J.min()) <= 1e-9 * J.min()
    for o in outs:
        # pipelined steps deliver the same winners one step late: the first returns the winner of the step launched
        # before the run (the last lock-step one), flush() the winner of the last
        assert o["pipelined_first"] == outs[0]["winners"][-1]
        assert o["pipelined"] == outs[0]["winners"]
        assert o["after_mix"] == outs[0]["winners"][0]
        assert o["host"] == outs[0]["winners"][:3]
    # explicit float32-stored controls through the host entry: job-wide winner of the 80,000-candidate job
    Ux = np.clip(np.random.default_rng(99).normal(size=(80_000, 12, 3)), -3, 3) * sigma
    Jx = orc.mpc_rollout_costs(p, Ux.astype(np.float32).astype(np.float64), pref)
    for o in outs:
        assert o["host_explicit"][0] == int(np.argmin(Jx))
        assert abs(o["host_explicit"][1] - Jx.min()) <= 1e-9 * Jx.min()
