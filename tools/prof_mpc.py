#!/usr/bin/env python
"""Minimal driver for ncu: a few launches of the config-3 MPC kernel (no CPU baseline, no e2e legs).
usage: prof_mpc.py [K] [H] [n_launch] [rng] [dtype]"""
import importlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import catenary_b200 as cb  # noqa: E402
import torch  # noqa: E402

K = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
H = int(sys.argv[2]) if len(sys.argv) > 2 else 50
N = int(sys.argv[3]) if len(sys.argv) > 3 else 10
rng = len(sys.argv) > 4 and sys.argv[4] == "rng"
dtype = np.float32 if len(sys.argv) > 5 and sys.argv[5] == "f32" else np.float64
dev = importlib.import_module(cb.__name__ + ".device")
sigma = cb.dynamics.load_model()["scaler_scale"][3:6]
p = cb.default_mpc_params()
pref = cb.circle_reference(H, p["dt"])
if rng:
    st = dev.MpcDeviceStep(p, pref, K, sigma=sigma, seed=1, dtype=dtype)
else:
    U = np.clip(np.random.default_rng(0).standard_normal((H, 3, K)), -3, 3) * np.asarray(sigma)[None, :, None]
    st = dev.MpcDeviceStep(p, pref, K, controls=U, dtype=dtype)
ts = []
for i in range(N):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    st.launch()
    b.record()
    b.synchronize()
    ts.append(a.elapsed_time(b))
print("K", K, "H", H, "rng", rng, dtype.__name__, "median ms", sorted(ts)[len(ts) // 2], "steps/s",
      K * H / (sorted(ts)[len(ts) // 2] * 1e-3), st.result())
