#!/usr/bin/env python
"""Back-to-back throughput of the config-3 rollout (the loop `value` times, with programmatic dependent launch):
200 launches between two events, explicit fp64 controls and device-sampled ones.  For tools/ab.sh."""
import importlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import catenary_b200 as cb  # noqa: E402
import torch  # noqa: E402

K, H, N = 65536, 50, 200
dev = importlib.import_module(cb.__name__ + ".device")
sigma = cb.dynamics.load_model()["scaler_scale"][3:6]
p = cb.default_mpc_params()
pref = cb.circle_reference(H, p["dt"])
sets = [torch.from_numpy(np.clip(np.random.default_rng(s).standard_normal((H, 3, K)), -3, 3) *
                         np.asarray(sigma)[None, :, None]).cuda() for s in range(3)]
out = []
for rng in (False, True):
    st = dev.MpcDeviceStep(p, pref, K, sigma=sigma, seed=1) if rng else dev.MpcDeviceStep(
        p, pref, K, controls=sets[0].cpu().numpy())
    for i in range(10):
        st.launch()
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(N):
            if not rng:
                st.set_controls(sets[i % 3])
            st.launch()
        b.record()
        b.synchronize()
        best = min(best, a.elapsed_time(b) / N)
    out.append(f"{'sampled' if rng else 'explicit'} {best * 1e3:.2f} us/step")
print("back-to-back:", ", ".join(out))
