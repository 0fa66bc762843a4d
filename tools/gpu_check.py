#!/usr/bin/env python
"""Developer smoke run on a GPU box: parity of every entry point against the oracle + rough timings.
(Not a test and not the bench; prints one line per check.)"""
import os
import sys
import time

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import catenary_b200 as cb                      # noqa: E402
from oracle import catenary_oracle as orc       # noqa: E402


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    m = np.isfinite(a) & np.isfinite(b)
    same_nan = np.array_equal(np.isnan(a), np.isnan(b))
    d = np.abs(a[m] - b[m]) / np.maximum(np.abs(b[m]), 1e-300)
    return (float(d.max()) if d.size else 0.0), same_nan


def main():
    ctx = cb.default_context()
    rng = np.random.default_rng(0)
    # solve_catenary
    n = 20000
    l = rng.uniform(0.05, 3.2, n)
    dH = rng.uniform(-2.5, 2.5, n)
    t0 = time.time()
    Cg = cb.solve_catenary(l, dH, 3.0)
    Co = orc.solve_catenary(l, dH, 3.0)
    print("solve_catenary rel", rel(Cg, Co), "nan frac", np.isnan(Co).mean())
    Cb = orc.solve_catenary_brentq(l[:300], dH[:300], 3.0)
    print("  vs brentq", rel(Cg[:300], Cb))
    Cf = cb.solve_catenary(l, dH, 3.0, dtype=np.float32)
    print("  f32", rel(Cf, Co))
    # shapes
    n = 3000
    A = rng.normal(size=(n, 3)) * 0.1
    dirs = rng.normal(size=(n, 3))
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    B = A + dirs * rng.uniform(0.3, 2.85, n)[:, None]
    th = rng.uniform(-0.5, 0.5, n)
    ga = rng.uniform(-0.5, 0.5, n)
    curves, Cs, valid = cb.catenary_shapes(A, B, 3.0, th, ga, which=15)
    worst = [0, 0, 0, 0]
    for i in range(0, n, 37):
        cat = orc.Catenary(3.0)
        ref = orc.transform_catenary(A[i], B[i], cat, th[i], ga[i])
        for j in range(4):
            if ref[j].shape[0] == 100:
                worst[j] = max(worst[j], np.abs(curves[j][i] - ref[j]).max())
    print("shapes abs err", worst, "valid hist", np.bincount(valid, minlength=4))
    # rk4 open loop
    T = 5000
    X = rng.normal(size=(T, 18))
    tt = np.arange(T) * 0.05
    for nm in ("dtheta_dt", "dgamma_dt"):
        yg = cb.rk4_integration(cb.EquationModel(nm), X, tt, 0.1)
        yo = orc.rk4_integration_vec(orc.EquationModel(nm), X, tt, 0.1)
        print("rk4", nm, np.abs(yg - yo).max())
    # all equations
    worst = 0
    for i, nm, cx, live in cb.equation_table():
        pg = cb.predict(X[:500], i)
        po = orc.equation(nm, cx)(X[:500])
        worst = max(worst, np.abs(pg - po).max() / max(1e-30, np.abs(po).max()))
    print("predict all eqs rel", worst)
    # MPC
    p = cb.default_mpc_params()
    sig = np.array(orc.load_model()["scaler_scale"][3:6])
    for K, H in ((4096, 50), (65536, 50)):
        U = np.clip(rng.normal(size=(K, H, 3)), -3, 3) * sig
        pref = cb.circle_reference(H, p["dt"])
        t0 = time.time()
        Jo = orc.mpc_rollout_costs(orc.default_mpc_params(), U, pref)
        t_or = time.time() - t0
        bi, bc, Jg = cb.mpc_sample_step(p, U, pref)
        io = int(np.argmin(Jo))
        print(f"mpc K={K} rel", rel(Jg, Jo), "argmin", bi, io, "cost", bc, Jo[io], f"oracle {t_or:.2f}s")
        Us = np.ascontiguousarray(U.transpose(1, 2, 0))
        bi2, bc2, Jg2 = cb.mpc_sample_step(p, Us, pref, layout="H3K")
        print("   SoA identical:", np.array_equal(Jg, Jg2), bi2 == bi)
        bi3, bc3, Jf = cb.mpc_sample_step(p, Us.astype(np.float32), pref, layout="H3K", dtype=np.float32)
        print("   f32 rel", rel(Jf, Jo), bi3)
        for _ in range(3):
            cb.mpc_sample_step(p, Us, pref, layout="H3K", return_costs=False)
        t0 = time.time()
        for _ in range(10):
            cb.mpc_sample_step(p, Us, pref, layout="H3K", return_costs=False)
        print(f"   host e2e {(time.time() - t0) / 10 * 1e3:.3f} ms/step")
    # RNG controls
    K, H = 8192, 50
    pref = cb.circle_reference(H, p["dt"])
    Uo = orc.sample_controls(1234, np.arange(K), H, sig)
    Jo = orc.mpc_rollout_costs(orc.default_mpc_params(), Uo, pref)
    bi, bc, Jg = cb.mpc_sample_step(p, None, pref, K=K, sigma=sig, seed=1234)
    print("mpc rng rel", rel(Jg, Jo), bi, int(np.argmin(Jo)))
    # generic equations
    pg = cb.default_mpc_params(theta_cx=19, gamma_cx=7)
    Jo = orc.mpc_rollout_costs(orc.default_mpc_params(theta_cx=19, gamma_cx=7), Uo[:2048], pref)
    _, _, Jg = cb.mpc_sample_step(pg, Uo[:2048], pref)
    print("mpc generic eq rel", rel(Jg, Jo))
    for kind in ("dfma", "ffma", "mufu"):
        print("peak", kind, f"{ctx.measure_peak(kind):.4e} ops/s")
    print("launches", ctx.launch_count)


if __name__ == "__main__":
    main()
