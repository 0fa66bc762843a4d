#!/usr/bin/env python
"""Device-resident timing of the config-2 shape kernel (cat_shape_dev_*): n configs x 100 points.
usage: prof_shape.py [n] [which] [dtype] [reps]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import catenary_b200 as cb  # noqa: E402
import torch  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
which = int(sys.argv[2]) if len(sys.argv) > 2 else 8
f32 = len(sys.argv) > 3 and sys.argv[3] == "f32"
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 10
dt = torch.float32 if f32 else torch.float64
rng = np.random.default_rng(0)
d = rng.normal(size=(n, 3))
d /= np.linalg.norm(d, axis=1, keepdims=True)
B = d * rng.uniform(1.2, 2.85, n)[:, None]
dev = torch.device("cuda", 0)
A_t = torch.zeros((n, 3), dtype=dt, device=dev)
B_t = torch.as_tensor(B, dtype=dt, device=dev)
L_t = torch.full((1,), 3.0, dtype=dt, device=dev)
th_t = torch.as_tensor(rng.uniform(-0.5, 0.5, n), dtype=dt, device=dev)
ga_t = torch.as_tensor(rng.uniform(-0.5, 0.5, n), dtype=dt, device=dev)
nout = bin(which).count("1")
out = torch.empty((nout, n, 100, 3), dtype=dt, device=dev)
C_t = torch.empty(n, dtype=dt, device=dev)
v_t = torch.empty(n, dtype=torch.uint8, device=dev)
ctx = cb.Context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
fn = ctx.lib.cat_shape_dev_f32 if f32 else ctx.lib.cat_shape_dev_f64
ts = []
for i in range(reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    ctx.check(fn(ctx.handle, n, A_t.data_ptr(), B_t.data_ptr(), L_t.data_ptr(), 0, th_t.data_ptr(), ga_t.data_ptr(), 100,
                 which, out.data_ptr(), C_t.data_ptr(), v_t.data_ptr()))
    b.record()
    b.synchronize()
    ts.append(a.elapsed_time(b))
ms = sorted(ts)[len(ts) // 2]
bytes_ = n * (72 + 9) * (4 if f32 else 8) / 8 + out.numel() * out.element_size()
print(f"n {n} which {which} {'f32' if f32 else 'f64'} median {ms:.3f} ms  {n / ms * 1e3:.3e} configs/s  "
      f"{bytes_ / ms / 1e6:.1f} GB/s algorithmic  valid {float((v_t & 1).float().mean()):.3f}")
