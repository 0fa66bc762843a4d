"""The reference's other feature layouts and its data ingest (rank 4) on the GPU.

``feature_bank`` computes every per-row term once (cat_feature_bank_f64: kinematics, np.gradient acceleration,
cable-relative surge/sway, the per-row catenary solve + tension the reference runs through
``np.vectorize(brentq)`` at ~89 us/row); the builders below keep the reference's names, arguments and column
orders and are column selections of that bank.  ``uniform_resample`` / ``load_and_resample_all`` keep
main_fun.py:128-160's surface with the interpolation on the device.  The smoothed / second-order layouts
(``build_*_features_valid``, ``compute_derivatives``, ``preprocess_signals``, ``features_dd``) run SciPy's
gaussian_filter1d / savgol_filter and np.gradient as device kernels (csrc/cat_kernels_feat.cuh).
"""
from __future__ import annotations

import numpy as np

from . import _lib
from .context import default_context

BANK_W = 32
BANK = {"P1": slice(0, 3), "V1": slice(3, 6), "A1": slice(6, 9), "unit_rel": slice(9, 12), "dist_clip": 12,
        "angle_proj": 13, "v_surge": 14, "v_sway": 15, "l": 16, "delta_H": 17, "C": 18, "T": 19,
        "angle_proj_tanh": 20, "a1_dot_unitrel": 21, "a1_perp_norm": 22, "v1_norm": 23, "a1_norm": 24,
        "acc_vel_ratio": 25, "theta": 26, "gamma": 27, "theta_prev": 28, "gamma_prev": 29, "dist": 30}

P0_COLS = ["rod_end X", "rod_end Y", "rod_end Z"]
P1_COLS = ["robot_cable_attach_point X", "robot_cable_attach_point Y", "robot_cable_attach_point Z"]
V1_COLS = ["rob_cor_speed X", "rob_cor_speed Y", "rob_cor_speed Z"]
TRAJECTORY_COLS = [f"br{r}_{n}" for r in (0, 1) for n in ("x", "y", "z", "phi", "theta", "psi", "u", "v", "w", "p",
                                                           "q", "r")]         # Rov_traj_gen.py:133-134


def experiment_columns(df):
    """The experimental CSV schema (simple.py:16-35) -> contiguous arrays in the units the reference uses:
    P0, P1 (T,3) in m (mm / 1000), V1 (T,3) in mm/s, time, theta, gamma (T,)."""
    g = lambda k: np.asarray(df[k], np.float64)
    stack = lambda ks: np.ascontiguousarray(np.stack([g(k) for k in ks], axis=1))
    return (stack(P0_COLS) / 1000, stack(P1_COLS) / 1000, stack(V1_COLS), np.ascontiguousarray(g("Time")),
            np.ascontiguousarray(g("Theta")), np.ascontiguousarray(g("Gamma")))


def feature_bank(df, L=3.0, cable_wet_weight=1.521, ctx=None):
    """(T, 32) term bank of an experiment DataFrame (or any mapping with the CSV's column names)."""
    ctx = ctx or default_context()
    P0, P1, V1, t, th, ga = experiment_columns(df)
    bank = np.empty((t.shape[0], BANK_W))
    ctx.check(ctx.lib.cat_feature_bank_f64(ctx.handle, t.shape[0], _lib.ptr(np.ascontiguousarray(P0)),
                                           _lib.ptr(np.ascontiguousarray(P1)), _lib.ptr(V1), _lib.ptr(t),
                                           _lib.ptr(th), _lib.ptr(ga), float(L), float(cable_wet_weight),
                                           _lib.ptr(bank)))
    return bank


def _cols(bank, names):
    parts = []
    for n in names:
        c = bank[:, BANK[n]]
        parts.append(c if c.ndim == 2 else c[:, None])
    return np.hstack(parts)


def build_theta_features(df, L, cable_wet_weight, ctx=None):
    """main_fun.py:267-340: [theta, v_surge, v_sway, l, delta_H, T, tanh(angle_proj), A1_y, |A1_perp|]."""
    b = feature_bank(df, L, cable_wet_weight, ctx)
    return np.hstack([_cols(b, ["theta", "v_surge", "v_sway", "l", "delta_H", "T", "angle_proj_tanh"]),
                      b[:, 7:8], _cols(b, ["a1_perp_norm"])])


def build_gamma_features(df, L, cable_wet_weight, ctx=None):
    """main_fun.py:342-415: [gamma, v_sway, v_surge, l, delta_H, T, tanh(angle_proj), A1_y, |A1_perp|]."""
    b = feature_bank(df, L, cable_wet_weight, ctx)
    return np.hstack([_cols(b, ["gamma", "v_sway", "v_surge", "l", "delta_H", "T", "angle_proj_tanh"]),
                      b[:, 7:8], _cols(b, ["a1_perp_norm"])])


def build_clean_features(df, ctx=None):
    """main_fun.py:197-265 (13 columns)."""
    b = feature_bank(df, ctx=ctx)
    v1n, a1n = b[:, 23:24], b[:, 24:25]
    return np.hstack([_cols(b, ["v_surge", "a1_dot_unitrel", "v_sway", "a1_perp_norm", "v1_norm", "a1_norm",
                                "acc_vel_ratio"]), v1n ** 2, a1n ** 2,
                      _cols(b, ["dist_clip", "angle_proj_tanh"]), b[:, 4:5], b[:, 7:8]])


def extract_features16(df, ctx=None):
    """main_fun.extract_features (main_fun.py:166-193): [P1, V1, A1, unit_rel, tension, angle_proj, theta, gamma]
    (the 18-column variant with the previous angles is dynamics.extract_features, simple.py:15-41)."""
    b = feature_bank(df, ctx=ctx)
    return _cols(b, ["P1", "V1", "A1", "unit_rel", "dist_clip", "angle_proj", "theta", "gamma"])


def resample_linear(time_in, values, time_out, ctx=None):
    """scipy interp1d(kind='linear', fill_value='extrapolate') of the columns of ``values`` (n_in, n_cols)."""
    ctx = ctx or default_context()
    ti = np.ascontiguousarray(time_in, np.float64)
    v = np.ascontiguousarray(values, np.float64)
    if v.ndim == 1:
        v = v[:, None]
    if ti.ndim != 1 or v.shape[0] != ti.shape[0] or ti.shape[0] < 2:
        raise ValueError("time_in (n_in >= 2,) and values (n_in, n_cols) expected")
    to = np.ascontiguousarray(time_out, np.float64)
    out = np.empty((to.shape[0], v.shape[1]))
    ctx.check(ctx.lib.cat_resample_linear_f64(ctx.handle, ti.shape[0], _lib.ptr(ti), v.shape[1], _lib.ptr(v),
                                              to.shape[0], _lib.ptr(to), _lib.ptr(out)))
    return out


UNIFORM_DT = 0.05      # main_fun.py:127


def uniform_resample(df, dt=UNIFORM_DT, ctx=None):
    """Drop-in for main_fun.uniform_resample (main_fun.py:128-145): a DataFrame whose 'Time' column is the grid
    np.arange(t[0], t[-1], dt) and whose other columns, in their original order, are linearly interpolated
    (and extrapolated) onto it -- all columns in ONE device pass (k_resample_linear)."""
    import pandas as pd
    if "Time" not in df.columns:
        raise ValueError("Missing 'Time' column in dataset.")            # the reference's message
    value_cols = df.columns.drop("Time")
    knots = df["Time"].to_numpy(dtype=np.float64)
    grid = np.arange(knots[0], knots[-1], dt)
    table = resample_linear(knots, df[value_cols].to_numpy(dtype=np.float64), grid, ctx)
    resampled = pd.DataFrame(table, columns=value_cols)
    resampled.insert(0, "Time", grid)
    return resampled


def load_and_resample_all(file_list, dt=UNIFORM_DT, ctx=None):
    """Drop-in for main_fun.load_and_resample_all (main_fun.py:148-159): every CSV that carries Theta, Gamma and Time
    is cleaned of rows missing any of them, resampled on its own grid and appended; others are skipped with a warning."""
    import warnings

    import pandas as pd
    needed = ["Theta", "Gamma", "Time"]
    frames = []
    for path in file_list:
        raw = pd.read_csv(path)
        if not set(needed) <= set(raw.columns):
            warnings.warn(f"{path}: no {'/'.join(c for c in needed if c not in raw.columns)} column, skipped",
                          stacklevel=2)
            continue
        frames.append(uniform_resample(raw.dropna(subset=needed), dt, ctx))
    return pd.concat(frames, ignore_index=True)


# ---- smoothed / second-order layouts (main_fun.py:504-655, 768-776, 811-871) -------------------------------------
def _gaussian_taps(sigma, truncate=4.0):
    """The weights scipy.ndimage.gaussian_filter1d(x, sigma) correlates with (order 0, truncate 4)."""
    radius = int(truncate * float(sigma) + 0.5)
    x = np.arange(-radius, radius + 1)
    phi = np.exp(-0.5 / (float(sigma) * float(sigma)) * x ** 2)
    return np.ascontiguousarray(phi / phi.sum())


def _savgol_tables(window, order):
    """Interior taps and the two edge-fit matrices of scipy.signal.savgol_filter(x, window, order), mode='interp'."""
    if window % 2 != 1 or order >= window:
        raise ValueError("window_length must be odd and larger than polyorder")
    h = window // 2
    centred = np.vander(np.arange(-h, h + 1, dtype=np.float64), order + 1, increasing=True)
    taps = np.linalg.pinv(centred)[0]
    pos = np.arange(window, dtype=np.float64)
    fit = np.linalg.pinv(np.vander(pos, order + 1, increasing=True))
    left = np.vander(pos[:h], order + 1, increasing=True) @ fit
    right = np.vander(pos[window - h:], order + 1, increasing=True) @ fit
    return (np.ascontiguousarray(taps), np.ascontiguousarray(left), np.ascontiguousarray(right))


def _is_uniform(t):
    """NumPy's own test for the uniform branch of np.gradient: every spacing bitwise equal."""
    d = np.diff(t)
    return int(d.size > 0 and bool((d == d[0]).all()))


def _signals(x):
    a = np.ascontiguousarray(x, np.float64)
    if a.ndim == 1:
        return a[None], True
    if a.ndim != 2:
        raise ValueError("a signal (len,) or a stack of signals (n, len) expected")
    return a, False


def gaussian_filter1d(x, sigma, ctx=None):
    """scipy.ndimage.gaussian_filter1d(x, sigma) along the last axis (mode 'reflect') on the GPU; x (len,) or (n, len)."""
    ctx = ctx or default_context()
    a, one = _signals(x)
    taps = _gaussian_taps(sigma)
    out = np.empty_like(a)
    ctx.check(ctx.lib.cat_fir_reflect_f64(ctx.handle, a.shape[0], a.shape[1], _lib.ptr(a), taps.size, _lib.ptr(taps),
                                          _lib.ptr(out)))
    return out[0] if one else out


def savgol_filter(x, window_length, polyorder, ctx=None):
    """scipy.signal.savgol_filter(x, window_length, polyorder) (mode 'interp') on the GPU; x (len,) or (n, len)."""
    ctx = ctx or default_context()
    a, one = _signals(x)
    if a.shape[1] < window_length:
        raise ValueError("If mode is 'interp', window_length must be less than or equal to the size of x.")
    taps, left, right = _savgol_tables(int(window_length), int(polyorder))
    out = np.empty_like(a)
    ctx.check(ctx.lib.cat_savgol_f64(ctx.handle, a.shape[0], a.shape[1], _lib.ptr(a), int(window_length), _lib.ptr(taps),
                                     _lib.ptr(left), _lib.ptr(right), _lib.ptr(out)))
    return out[0] if one else out


def gradient(y, t, ctx=None):
    """np.gradient(y, t) (edge_order 1; uniform and non-uniform branches) on the GPU; y (len,) or (n, len)."""
    ctx = ctx or default_context()
    a, one = _signals(y)
    tt = np.ascontiguousarray(t, np.float64)
    if tt.shape != (a.shape[1],) or tt.size < 2:
        raise ValueError("t must have the signal's length (>= 2)")
    out = np.empty_like(a)
    ctx.check(ctx.lib.cat_gradient_f64(ctx.handle, a.shape[0], a.shape[1], _lib.ptr(a), _lib.ptr(tt), _is_uniform(tt),
                                       _lib.ptr(out)))
    return out[0] if one else out


def preprocess_signals(df, sigma=2, ctx=None):
    """main_fun.preprocess_signals (main_fun.py:768-776): (time, Gaussian-smoothed Theta, Gaussian-smoothed Gamma)."""
    t = np.ascontiguousarray(np.asarray(df["Time"], np.float64))
    sm = gaussian_filter1d(np.stack([np.asarray(df["Theta"], np.float64), np.asarray(df["Gamma"], np.float64)]),
                           sigma, ctx)
    return t, sm[0], sm[1]


def compute_derivatives(df, ctx=None):
    """main_fun.compute_derivatives (main_fun.py:645-655): second time derivatives of the Savitzky-Golay (11, 3)
    smoothed Theta and Gamma -> (ddtheta, ddgamma)."""
    t = np.asarray(df["Time"], np.float64)
    sm = savgol_filter(np.stack([np.asarray(df["Theta"], np.float64), np.asarray(df["Gamma"], np.float64)]), 11, 3, ctx)
    dd = gradient(gradient(sm, t, ctx), t, ctx)
    return dd[0], dd[1]


def _valid(df, for_gamma, ctx):
    ctx = ctx or default_context()
    P0, P1, V1, t, th, ga = experiment_columns(df)
    taps = _gaussian_taps(2)                                             # sigma = 2, main_fun.py:510-518
    out = np.empty((t.shape[0], 13))
    ctx.check(ctx.lib.cat_features_valid_f64(ctx.handle, t.shape[0], _lib.ptr(P0), _lib.ptr(P1), _lib.ptr(V1),
                                             _lib.ptr(t), _lib.ptr(th), _lib.ptr(ga), taps.size, _lib.ptr(taps),
                                             _is_uniform(t), int(for_gamma), _lib.ptr(out)))
    return out


def build_theta_features_valid(df, L=3.0, cable_wet_weight=1.521, ctx=None):
    """main_fun.build_theta_features_valid (main_fun.py:504-571), 13 columns: theta v_surge / l, v_surge / l,
    theta v_surge, v_surge, v_surge^2 / l, theta, gamma, gamma v_sway / l, v_sway / l, gamma v_sway, v_sway, dtheta,
    dgamma.  (L and cable_wet_weight are accepted as in the reference, whose tension term is computed but not stacked.)"""
    return _valid(df, 0, ctx)


def build_gamma_features_valid(df, L=3.0, cable_wet_weight=1.521, ctx=None):
    """main_fun.build_gamma_features_valid (main_fun.py:574-642): the sway / gamma terms first."""
    return _valid(df, 1, ctx)


def features_dd(df, ctx=None):
    """main_fun.features_dd (main_fun.py:811-871) -> (features (T,14), targets (T,2)): the layout the second-order
    (theta_ddot, gamma_ddot) models of dd_test_cluster.py / test_cluster.py:110-129 are trained on."""
    ctx = ctx or default_context()
    P0, P1, V1, t, th, ga = experiment_columns(df)
    if t.shape[0] < 11:
        raise ValueError("If mode is 'interp', window_length must be less than or equal to the size of x.")
    taps, left, right = _savgol_tables(11, 3)                            # main_fun.py:829-830
    feats = np.empty((t.shape[0], 14))
    targets = np.empty((t.shape[0], 2))
    ctx.check(ctx.lib.cat_features_dd_f64(ctx.handle, t.shape[0], _lib.ptr(P0), _lib.ptr(P1), _lib.ptr(V1), _lib.ptr(t),
                                          _lib.ptr(th), _lib.ptr(ga), 11, _lib.ptr(taps), _lib.ptr(left),
                                          _lib.ptr(right), _is_uniform(t), _lib.ptr(feats), _lib.ptr(targets)))
    return feats, targets


def load_trajectory_csv(path):
    """The 24-column trajectory CSV of Rov_traj_gen.py:133-139 -> (2, T, 12) float64: per vehicle
    [x, y, z, phi, theta, psi, u, v, w, p, q, r]."""
    a = np.loadtxt(path, delimiter=",", skiprows=1, ndmin=2)
    if a.shape[1] != 24:
        raise ValueError(f"{path}: expected 24 columns, found {a.shape[1]}")
    return np.ascontiguousarray(a.reshape(a.shape[0], 2, 12).transpose(1, 0, 2))


def config1_trajectory(T=5000, seed=42):
    """deterministic synthetic experiment in the CSV's data units -- P1 on a 0.4 m circle
    (period 10 s) with a 0.25 m heave (period 25 s), V1 = dP1/dt + a +-30 mm/s PRBS on x (default_rng(42),
    the seed of simple.py:77), Theta = 0.15 sin(2 pi t / 7), Gamma = 0.10 cos(2 pi t / 9), dt = 0.05 s."""
    t = np.arange(T) * 0.05
    rng = np.random.default_rng(seed)
    w10, w25 = 2 * np.pi / 10, 2 * np.pi / 25
    P1 = 1000 * np.stack([1 + 0.4 * np.cos(w10 * t), 1 + 0.4 * np.sin(w10 * t), -0.5 + 0.25 * np.sin(w25 * t)], 1)
    V1 = 1000 * np.stack([-0.4 * w10 * np.sin(w10 * t), 0.4 * w10 * np.cos(w10 * t), 0.25 * w25 * np.cos(w25 * t)], 1)
    V1[:, 0] += 30.0 * (2 * rng.integers(0, 2, T) - 1)
    cols = {"Time": t, "Theta": 0.15 * np.sin(2 * np.pi * t / 7), "Gamma": 0.10 * np.cos(2 * np.pi * t / 9)}
    for j, ax in enumerate("XYZ"):
        cols[f"rod_end {ax}"] = np.zeros(T)
        cols[f"robot_cable_attach_point {ax}"] = P1[:, j]
        cols[f"rob_cor_speed {ax}"] = V1[:, j]
    return cols


def config1_pipeline(cols, L=3.0, ctx=None):
    """The reference's config-1 flow on the GPU: extract_features + scaler (simple.py:15-41, test_temp.py:49) ->
    rk4_integration of both ODEs (simulate_rk4_theta_gamma.py:74-75) -> per-step solve_catenary
    (main_fun.py:418-431) -> transform_catenary with the integrated angles (fully_augmented_catenary.py:18).
    -> dict(theta (T,), gamma (T,), C (T,), curves (T,100,3), valid (T,))."""
    from .dynamics import EquationModel, extract_features, rk4_integration
    from .geometry import catenary_shapes
    X = extract_features(cols, scaled=True, ctx=ctx)
    t = np.asarray(cols["Time"], np.float64)
    theta = rk4_integration(EquationModel("dtheta_dt", ctx=ctx), X, t, float(cols["Theta"][0]), ctx=ctx)
    gamma = rk4_integration(EquationModel("dgamma_dt", ctx=ctx), X, t, float(cols["Gamma"][0]), ctx=ctx)
    P0, P1, _, _, _, _ = experiment_columns(cols)
    (curves,), C, valid = catenary_shapes(P0, P1, L, theta, gamma, ctx=ctx)
    return dict(theta=theta, gamma=gamma, C=C, curves=curves, valid=valid)
