"""Oracle restatements of the reference's other feature layouts / resampler against vectors produced by the
reference's own main_fun functions (oracle/make_golden_features.py)."""
import numpy as np

from conftest import assert_close, golden
from oracle import catenary_oracle as orc

COLS = ["Time"] + [f"{p} {a}" for p in ("rod_end", "robot_cable_attach_point", "rob_cor_speed") for a in "XYZ"] + [
    "Theta", "Gamma"]


def frame(g):
    return {c: g[c] for c in COLS}


def test_feature_layouts_match_reference():
    g = golden("feature_layouts.npz")
    df = frame(g)
    # tension goes through solve_catenary: the oracle's root is closer to the true root than brentq's
    # (xtol 2e-12), hence 1e-8 on that column only
    for name, fn, args in (("theta", orc.build_theta_features, (3.0, 1.521)),
                           ("gamma", orc.build_gamma_features, (3.0, 1.521))):
        got, want = fn(df, *args), g[f"ref_{name}_features"]
        keep = [j for j in range(9) if j != 5]
        assert_close(got[:, keep], want[:, keep], 1e-12, 1e-12, name)
        assert_close(got[:, 5], want[:, 5], 1e-8, 0, name + " tension")
    assert_close(orc.build_clean_features(df), g["ref_clean_features"], 1e-12, 1e-12, "clean")
    assert_close(orc.extract_features16(df), g["ref_features16"], 1e-12, 1e-12, "features16")
    assert np.isnan(orc.feature_terms(df)["C"]).sum() >= 15        # the taut / near-vertical rows are in the set


def test_resample_matches_reference():
    g = golden("feature_layouts.npz")
    names = list(g["ref_resampled_columns"])
    assert names[0] == "Time"
    others = names[1:]
    vals = np.stack([g[c] for c in others], 1)
    for key, dt in (("ref_resampled", 0.05), ("ref_resampled_fine", 0.0125)):
        t_new = np.arange(g["Time"][0], g["Time"][-1], dt)
        want = g[key]
        assert np.array_equal(want[:, 0], t_new)
        got = orc.resample_linear(g["Time"], vals, t_new)
        assert np.array_equal(got, want[:, 1:]), key             # same formula as scipy: bit-exact


def test_smoothed_and_second_order_layouts_match_reference():
    """build_*_features_valid, compute_derivatives, preprocess_signals, features_dd (main_fun.py:504-655, 768-776,
    811-871) restated in NumPy -- gaussian_filter1d's weights and 'reflect' boundary, savgol_filter's interior taps and
    polynomial edge fit, np.gradient's uniform and non-uniform branches -- against the reference's own outputs with
    the real SciPy, on a jittered and on an exactly uniform time grid."""
    g, g2 = golden("feature_layouts.npz"), golden("feature_layouts2.npz")
    for tag in ("", "u_"):
        df = frame(g)
        df["Time"] = g2[tag + "Time"]
        assert_close(orc.build_theta_features_valid(df), g2[tag + "ref_theta_features_valid"], 1e-10, 1e-12, tag + "theta valid")
        assert_close(orc.build_gamma_features_valid(df), g2[tag + "ref_gamma_features_valid"], 1e-10, 1e-12, tag + "gamma valid")
        ddt, ddg = orc.compute_derivatives(df)
        assert_close(ddt, g2[tag + "ref_ddtheta"], 1e-9, 1e-9, tag + "ddtheta")   # second differences of 4e-3 noise: |dd| ~ 10
        assert_close(ddg, g2[tag + "ref_ddgamma"], 1e-9, 1e-9, tag + "ddgamma")
        _, ths, gas = orc.preprocess_signals(df)
        assert_close(ths, g2[tag + "ref_pre_theta"], 1e-12, 1e-14, tag + "gaussian theta")
        assert_close(gas, g2[tag + "ref_pre_gamma"], 1e-12, 1e-14, tag + "gaussian gamma")
        _, ths, gas = orc.preprocess_signals(df, sigma=3.5)
        assert_close(ths, g2[tag + "ref_pre_theta_s35"], 1e-12, 1e-14, tag + "gaussian theta sigma 3.5")
        f, tg = orc.features_dd(df)
        assert_close(f, g2[tag + "ref_dd_features"], 1e-9, 1e-10, tag + "features_dd")
        assert_close(tg, g2[tag + "ref_dd_targets"], 1e-9, 1e-9, tag + "features_dd targets")
    taps, el, er = orc.savgol_tables(11, 3)
    assert abs(taps.sum() - 1) < 1e-14 and np.allclose(taps, taps[::-1]) and np.allclose(el.sum(1), 1) and np.allclose(er.sum(1), 1)
