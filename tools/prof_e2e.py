#!/usr/bin/env python
"""Host MPC step (mpc_sample_step, config 3) from page-locked candidates: median ms per call, float32 and fp64 on the
wire.  Environment switches of the library apply (CAT_MPC_STREAMED, CAT_MPC_STREAM_GROUPS, CAT_MPC_STREAM_MEMOP).
usage: prof_e2e.py [reps=40]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import catenary_b200 as cb  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
K, H = 65536, 50
ctx = cb.default_context()
p = cb.default_mpc_params()
pref = cb.circle_reference(H, p["dt"])
rng = np.random.default_rng(1)
sig = np.array([108.5, 15.9, 63.1])
out = []
for dt in (np.float32, np.float64):
    sets = []
    for i in range(4):
        a = ctx.pinned_empty((H, 3, K), dt)
        a[:] = np.clip(rng.normal(size=(H, 3, K)), -3, 3) * sig[None, :, None]
        sets.append(a)
    kw = dict(layout="H3K", return_costs=False, ctx=ctx)
    if dt is np.float32:
        kw["controls_dtype"] = np.float32
    for i in range(5):
        cb.mpc_sample_step(p, sets[i % 4], pref, **kw)
    ts = []
    for i in range(reps):
        t0 = time.perf_counter()
        cb.mpc_sample_step(p, sets[i % 4], pref, **kw)
        ts.append(time.perf_counter() - t0)
    out.append(f"{np.dtype(dt).name} median {np.median(ts) * 1e3:.4f} ms min {np.min(ts) * 1e3:.4f}")
print(" | ".join(out), "|", {k: v for k, v in os.environ.items() if k.startswith("CAT_MPC_")})
