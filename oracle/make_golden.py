#!/usr/bin/env python
"""Generate tests/golden/*.npz by EXECUTING the reference's own functions (build container only).

TEST INFRASTRUCTURE.  Run:  python oracle/make_golden.py
Every array named ``ref_*`` is the output of unmodified reference code (oracle/reference_loader.py) on the
inputs stored beside it; nothing here calls the oracle restatement except ``Catenary`` (the stand-in for the
missing pympc class, used only as the ``catenary_fn`` ARGUMENT of the reference's transform_catenary) and
``EquationModel`` (the lambdified CSV equation used as the ``model`` ARGUMENT of the reference's integrators,
standing in for the un-loadable PySR pickle).
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "..")
sys.path.insert(0, ROOT)
from oracle import catenary_oracle as orc          # noqa: E402
from oracle import reference_loader as ref         # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def synthetic_trajectory(T, seed=42):
    """Config-1 trajectory in data units: columns of the experimental CSV schema."""
    t = np.arange(T) * 0.05
    rng = np.random.default_rng(seed)
    w10, w25 = 2 * np.pi / 10, 2 * np.pi / 25
    P1 = 1000 * np.stack([1 + 0.4 * np.cos(w10 * t), 1 + 0.4 * np.sin(w10 * t), -0.5 + 0.25 * np.sin(w25 * t)], 1)
    V1 = 1000 * np.stack([-0.4 * w10 * np.sin(w10 * t), 0.4 * w10 * np.cos(w10 * t), 0.25 * w25 * np.cos(w25 * t)], 1)
    V1[:, 0] += 30.0 * (2 * rng.integers(0, 2, T) - 1)
    cols = {"Time": t, "rod_end X": np.zeros(T), "rod_end Y": np.zeros(T), "rod_end Z": np.zeros(T),
            "robot_cable_attach_point X": P1[:, 0], "robot_cable_attach_point Y": P1[:, 1],
            "robot_cable_attach_point Z": P1[:, 2], "rob_cor_speed X": V1[:, 0], "rob_cor_speed Y": V1[:, 1],
            "rob_cor_speed Z": V1[:, 2], "Theta": 0.15 * np.sin(2 * np.pi * t / 7),
            "Gamma": 0.10 * np.cos(2 * np.pi * t / 9)}
    return cols


def main():
    import pandas as pd
    os.makedirs(OUT, exist_ok=True)
    mf = ref.main_fun()
    rng = np.random.default_rng(2025)

    # ---- rodrigues_rotation, main_fun.py:18-35 ----
    v = rng.normal(size=(64, 3))
    ax = rng.normal(size=(64, 3)) * rng.uniform(0.1, 5, (64, 1))
    an = rng.uniform(-np.pi, np.pi, 64)
    out = np.array([mf.rodrigues_rotation(v[i], ax[i], an[i]) for i in range(64)])
    np.savez(os.path.join(OUT, "rodrigues.npz"), v=v, axis=ax, angle=an, ref_out=out)

    # ---- solve_catenary, main_fun.py:418-431 (scipy brentq; nan domain) ----
    l = np.concatenate([[1.0, np.sqrt(2.0), 2.0, 2.9], rng.uniform(0.02, 3.3, 596)])
    dH = np.concatenate([[0.0, -1.0, 0.5, 0.0], rng.uniform(-2.9, 2.9, 596)])
    with np.errstate(all="ignore"):
        C3 = mf.solve_catenary(l, dH, 3.0)
        L2 = rng.uniform(2.0, 5.0, 600)
        CL = np.array([float(mf.solve_catenary(l[i], dH[i], L2[i])) for i in range(600)])
    np.savez(os.path.join(OUT, "solve_catenary.npz"), l=l, dH=dH, ref_C_L3=np.asarray(C3, np.float64), L=L2, ref_C=CL)

    # ---- transform_catenary, main_fun.py:38-111, catenary_fn = stand-in Catenary ----
    scen = [  # (A, B, L, theta, gamma)  literal scenarios of the reference scripts + edge cases
        ((0, 0, 0), (1, 1, 0), 3.0, 0.0, 0.0),                                   # catenary.py:10-14
        ((0, 0, 0), (1, 1, -1), 3.0, np.radians(-10), np.radians(20)),           # fully_augmented_catenary.py:11-15
        ((0, 0, 0), (1, 1, -1), 3.0, np.radians(10), np.radians(15)),            # Catenary_movement_theta_gama.py:13-14
        ((0, 0, 0), (1, 1, -1), 3.0, np.radians(-30), np.radians(20)),           # Animation.py:55,86
        ((0.2, -0.1, 0.3), (1.5, 0.7, -0.8), 3.0, 0.3, -0.4),
        ((0, 0, 0), (0, 0, -2.0), 3.0, 0.2, 0.1),                                # vertical: axis fallback, nan C
        ((0, 0, 0), (2.5, 2.0, 0.5), 3.0, 0.1, 0.2),                             # taut: None -> [A, B]
        ((0, 0, 10), (10, 0, 10), 12.0, 0.0, 0.5),                               # models/catenary_3d.py:49-52
    ]
    for _ in range(24):
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        A = rng.normal(size=3) * 0.2
        scen.append((tuple(A), tuple(A + d * rng.uniform(0.3, 2.85)), 3.0, rng.uniform(-0.5, 0.5), rng.uniform(-0.5, 0.5)))
    A = np.array([s[0] for s in scen], np.float64)
    B = np.array([s[1] for s in scen], np.float64)
    Ls = np.array([s[2] for s in scen])
    th = np.array([s[3] for s in scen])
    ga = np.array([s[4] for s in scen])
    npts = np.zeros((len(scen), 2), np.int64)
    curves = np.full((len(scen), 4, 100, 3), np.nan)
    for i in range(len(scen)):
        res = mf.transform_catenary(A[i], B[i], orc.Catenary(Ls[i], "ENU"), th[i], ga[i])
        npts[i] = (res[0].shape[0], res[1].shape[0])
        for j in range(4):
            curves[i, j, :res[j].shape[0]] = res[j]
    np.savez(os.path.join(OUT, "transform_catenary.npz"), A=A, B=B, L=Ls, theta=th, gamma=ga, n_pts=npts,
             ref_curves=curves)

    # ---- extract_features (18 col), simple.py:15-41; StandardScaler; integrators ----
    cols = synthetic_trajectory(400)
    df = pd.DataFrame(cols)
    X = ref.extract_features18()(df)
    sc = ref.scaler()
    Xs = sc.transform(X)
    rk4 = ref.rk4_integration()
    t = cols["Time"]
    # non-uniform time grid for the integrators (dt = time[i]-time[i-1], simulate_rk4_theta_gamma.py:55)
    t_nu = np.cumsum(rng.uniform(0.02, 0.08, 400))
    res = {}
    for nm in ("dtheta_dt", "dgamma_dt"):
        m = orc.EquationModel(nm)
        res[f"ref_rk4_{nm}"] = rk4(m, Xs, t, 0.05)
        res[f"ref_rk4_nu_{nm}"] = rk4(m, Xs, t_nu, -0.02)
        res[f"ref_predict_{nm}"] = m.predict(Xs)
    th_e, ga_e = mf.integrate_theta_gamma(orc.EquationModel("dtheta_dt"), orc.EquationModel("dgamma_dt"), Xs, t,
                                          0.05, 0.1)
    np.savez(os.path.join(OUT, "dynamics.npz"), **{k: np.asarray(v) for k, v in cols.items()}, ref_X=X, ref_Xs=Xs,
             time_nu=t_nu, ref_euler_theta=th_e, ref_euler_gamma=ga_e, **res)

    # ---- compute_catenary_3D, models/catenary_3d.py:5-39 ----
    c3d = ref.compute_catenary_3D()
    p0 = np.array([[0, 0, 10.0], [0, 0, 0], [0.1, 0.2, 0.3], [0, 0, 0]])
    p1 = np.array([[10, 0, 10.0], [1, 1, -1], [1.4, -0.6, 0.9], [2.5, 2.0, 0.5]])
    LL = np.array([12.0, 3.0, 3.0, 3.0])
    pts = np.array([c3d(p0[i], p1[i], LL[i], 100) for i in range(4)])
    np.savez(os.path.join(OUT, "catenary_3d.npz"), p0=p0, p1=p1, L=LL, ref_pts=pts)
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
