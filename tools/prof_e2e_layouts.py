#!/usr/bin/env python
"""Host MPC step (config 3) by candidate layout and wire type: (K,H,3) = the oracle's AoS (chunked upload) against
(H,3,K) = the kernel's SoA (streamed upload).  One B200, session 9: float32 0.821 / 0.774 ms, float64 1.533 / 1.482 ms."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import catenary_b200 as cb
K, H = 65536, 50
ctx = cb.default_context()
p = cb.default_mpc_params()
pref = cb.circle_reference(H, p["dt"])
rng = np.random.default_rng(1)
sig = np.array([108.5, 15.9, 63.1])
for dt in (np.float32, np.float64):
    for layout, shape in (("KH3", (K, H, 3)), ("H3K", (H, 3, K))):
        sets = []
        for i in range(4):
            a = ctx.pinned_empty(shape, dt)
            z = np.clip(rng.normal(size=(K, H, 3)), -3, 3) * sig
            a[:] = z if layout == "KH3" else z.transpose(1, 2, 0)
            sets.append(a)
        kw = dict(layout=layout, return_costs=False, ctx=ctx)
        if dt is np.float32:
            kw["controls_dtype"] = np.float32
        for i in range(5):
            r = cb.mpc_sample_step(p, sets[i % 4], pref, **kw)
        ts = []
        for i in range(40):
            t0 = time.perf_counter()
            cb.mpc_sample_step(p, sets[i % 4], pref, **kw)
            ts.append(time.perf_counter() - t0)
        print(np.dtype(dt).name, layout, f"median {np.median(ts)*1e3:.4f} ms", r[:2])
