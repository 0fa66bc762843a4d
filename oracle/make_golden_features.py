#!/usr/bin/env python
"""Golden vectors for the other feature layouts and the resampler (TEST INFRASTRUCTURE; needs /root/reference).

Imports the reference's main_fun UNMODIFIED (oracle/reference_loader.py) and runs its own functions on a seeded
synthetic experiment DataFrame: build_theta_features / build_gamma_features (main_fun.py:267-415, incl. the
np.vectorize(brentq) solve_catenary per row), build_clean_features (main_fun.py:197-265), extract_features
(main_fun.py:167-193), uniform_resample (main_fun.py:128-145).  Writes tests/golden/feature_layouts.npz, and -- for the
smoothed / second-order layouts build_theta_features_valid, build_gamma_features_valid (main_fun.py:504-642),
compute_derivatives (:645-655), preprocess_signals (:768-776), features_dd (:811-871), on the jittered and on an exactly
uniform time grid -- tests/golden/feature_layouts2.npz (inputs are those of feature_layouts.npz)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
from oracle import reference_loader as ref  # noqa: E402

OUT = os.path.join(HERE, "..", "tests", "golden")


def synthetic_experiment(T=600, seed=7):
    """Experimental-CSV schema (simple.py:16-35), data units (mm, mm/s, rad, s); jittered time stamps; a moving
    anchor; a stretch of rows beyond the cable length (taut: solve_catenary -> nan -> tension fallback) and a
    few near-vertical spans."""
    rng = np.random.default_rng(seed)
    t = np.cumsum(rng.uniform(0.03, 0.07, T))
    w1, w2 = 2 * np.pi / 11, 2 * np.pi / 17
    P0 = 1000 * np.stack([0.05 * np.sin(w2 * t), 0.05 * np.cos(w2 * t), 0.02 * np.sin(w1 * t)], 1)
    r = 1.6 + 0.7 * np.sin(w2 * t)
    P1 = 1000 * np.stack([r * np.cos(w1 * t), r * np.sin(w1 * t), -0.6 + 0.4 * np.sin(w2 * t + 1)], 1)
    P1[200:215] *= (3300.0 / np.linalg.norm(P1[200:215], axis=1))[:, None]   # |P1 - P0| ~ 3.3 m > L = 3 m
    P1[300:304, :2] = P0[300:304, :2] + 1e-4                # near-vertical
    V1 = np.gradient(P1, t, axis=0) + rng.normal(size=(T, 3)) * 15.0
    V1[400:403] = 0.0                                       # zero velocity: the +1e-8 guards
    cols = {"Time": t}
    for j, ax in enumerate("XYZ"):
        cols[f"rod_end {ax}"] = P0[:, j]
        cols[f"robot_cable_attach_point {ax}"] = P1[:, j]
        cols[f"rob_cor_speed {ax}"] = V1[:, j]
    cols["Theta"] = 0.2 * np.sin(2 * np.pi * t / 7) + rng.normal(size=T) * 0.004
    cols["Gamma"] = 0.12 * np.cos(2 * np.pi * t / 9) + rng.normal(size=T) * 0.004
    return cols


def uniform_time(cols, dt=0.0625):
    """The same experiment on an EXACTLY uniform time grid (dt a power of two): np.gradient then takes its uniform
    branch ((f[i+1] - f[i-1]) / 2h) instead of the non-uniform second-order formula."""
    c = dict(cols)
    c["Time"] = np.arange(len(cols["Time"])) * dt
    return c


def main():
    import pandas as pd
    mf = ref.main_fun()
    cols = synthetic_experiment()
    df = pd.DataFrame(cols)
    L, w = 3.0, 1.521                                       # catenary.py:10, test_cluster.py:22-23
    out = {k: np.asarray(v) for k, v in cols.items()}
    out["ref_theta_features"] = mf.build_theta_features(df, L, w)
    out["ref_gamma_features"] = mf.build_gamma_features(df, L, w)
    out["ref_clean_features"] = mf.build_clean_features(df)
    out["ref_features16"] = mf.extract_features(df)
    rs = mf.uniform_resample(df, 0.05)
    out["ref_resampled_columns"] = np.array(list(rs.columns))
    out["ref_resampled"] = rs.values
    rs2 = mf.uniform_resample(df, 0.0125)
    out["ref_resampled_fine"] = rs2.values
    np.savez_compressed(os.path.join(OUT, "feature_layouts.npz"), **out)
    # ---- catenary_model.py:10-31 (level-support planar profile), lifted defs, scipy's bounded minimiser ----
    from scipy.optimize import minimize_scalar
    cm = ref.lift(os.path.join("models", "catenary-3d-visualization", "src", "catenary_model.py"),
                  ["catenary_equation", "solve_catenary_params", "generate_catenary"],
                  {"minimize_scalar": minimize_scalar})
    cases = np.array([[3.0, 1.0], [3.0, 2.0], [3.0, 2.9], [12.0, 10.0], [5.0, 0.5], [3.0, 2.999]])
    Cs, xs, ys = [], [], []
    for L_, l_ in cases:
        Cs.append(cm["solve_catenary_params"](L_, l_, 0.0))
        x, y = cm["generate_catenary"](None, L_, l_, 0.0)
        xs.append(x)
        ys.append(y)
    xe = np.linspace(-2, 2, 41)
    np.savez_compressed(os.path.join(OUT, "catenary_model.npz"), cases=cases, ref_C=np.array(Cs), ref_x=np.array(xs),
                        ref_y=np.array(ys), eq_x=xe, eq_C=np.array([0.05, 0.7, 3.0]),
                        ref_eq=np.array([cm["catenary_equation"](xe, c) for c in (0.05, 0.7, 3.0)]))
    # ---- second-order / smoothed layouts (second half): main_fun.py:504-655, 768-776, 811-871 ----
    out2 = {}
    for tag, cols_t in (("", cols), ("u_", uniform_time(cols))):
        dft = pd.DataFrame(cols_t)
        out2[tag + "Time"] = np.asarray(cols_t["Time"])
        out2[tag + "ref_theta_features_valid"] = mf.build_theta_features_valid(dft, L, w)
        out2[tag + "ref_gamma_features_valid"] = mf.build_gamma_features_valid(dft, L, w)
        dd_t, dd_g = mf.compute_derivatives(dft)
        out2[tag + "ref_ddtheta"], out2[tag + "ref_ddgamma"] = dd_t, dd_g
        tt, th_s, ga_s = mf.preprocess_signals(dft, sigma=2)
        out2[tag + "ref_pre_theta"], out2[tag + "ref_pre_gamma"] = th_s, ga_s
        _, th_s3, ga_s3 = mf.preprocess_signals(dft, sigma=3.5)
        out2[tag + "ref_pre_theta_s35"], out2[tag + "ref_pre_gamma_s35"] = th_s3, ga_s3
        f_dd, t_dd = mf.features_dd(dft)
        out2[tag + "ref_dd_features"], out2[tag + "ref_dd_targets"] = f_dd, t_dd
    np.savez_compressed(os.path.join(OUT, "feature_layouts2.npz"), **out2)
    print({k: v.shape for k, v in out2.items() if "ref_" in k})
    print({k: v.shape for k, v in out.items() if k.startswith("ref_")},
          "nan tension rows:", int(np.isnan(mf.solve_catenary(out["ref_theta_features"][:, 3],
                                                              out["ref_theta_features"][:, 4], L)).sum()))


if __name__ == "__main__":
    main()
