"""Fused peer-memory winner exchange on ONE GPU: two ranks (gloo for the handle swap) share
cuda:0, map each other's slot buffers through CUDA IPC and run the one-kernel step.  The two processes' kernels are
time-sliced, so this checks the protocol (sequence numbers, parities, rank-order fold), not the latency."""
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
out = {"rank": rank, "winners": []}
try:
    for seed in (3, 4, 5, 6, 7):          # five steps: both parities reused, a fresh job each time
        st = devmod.MpcDeviceStep(p, pref, hi - lo, sigma=sigma, seed=seed, k_offset=lo, device=0)
        if seed == 3:
            ex = dmod.P2PWinnerExchange(st)
        else:
            ex.step = st
        ex.launch()
        out["winners"].append(list(ex.result()))
    out["timed_out"] = ex.timed_out()
except Exception as e:                    # IPC refused in this environment, ...
    out["error"] = repr(e)
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
    for seed, (bi, bc) in zip((3, 4, 5, 6, 7), outs[0]["winners"]):
        J = orc.mpc_rollout_costs(p, orc.sample_controls(seed, np.arange(6000), 12, sigma), pref)
        assert bi == int(np.argmin(J)) and abs(bc - J.min()) <= 1e-9 * J.min()
