"""The reference's other feature layouts, tension and resampler on the GPU against the
vectors its own main_fun functions produced (tests/golden/feature_layouts.npz) and against the oracle."""
import numpy as np
import pytest

from conftest import assert_close, golden
from oracle import catenary_oracle as orc
from test_features_oracle_cpu import frame

pytestmark = pytest.mark.gpu


def test_feature_layouts_match_reference(cb):
    g = golden("feature_layouts.npz")
    df = frame(g)
    keep = [j for j in range(9) if j != 5]
    for name, fn in (("theta", cb.build_theta_features), ("gamma", cb.build_gamma_features)):
        got, want = fn(df, 3.0, 1.521), g[f"ref_{name}_features"]
        assert_close(got[:, keep], want[:, keep], 1e-9, 1e-12, name)
        assert_close(got[:, 5], want[:, 5], 1e-8, 0, name + " tension (reference root = brentq, xtol 2e-12)")
        assert_close(got, getattr(orc, f"build_{name}_features")(df, 3.0, 1.521), 1e-9, 1e-12, name + " vs oracle")
    assert_close(cb.build_clean_features(df), g["ref_clean_features"], 1e-9, 1e-12, "clean")
    assert_close(cb.extract_features16(df), g["ref_features16"], 1e-9, 1e-12, "features16")
    bank = cb.feature_bank(df)
    f = orc.feature_terms(df)
    assert_close(bank[:, 18], f["C"][:, 0], 1e-9, 0, "C (NaN pattern included)")
    assert np.isnan(bank[:, 18]).sum() >= 15
    assert_close(bank[:, 28], np.r_[g["Theta"][0], g["Theta"][:-1]], 0, 0, "theta_prev")


def test_feature_bank_pandas_and_large(cb):
    pd = pytest.importorskip("pandas")
    rng = np.random.default_rng(12)
    T = 200_000
    t = np.cumsum(rng.uniform(0.04, 0.06, T))
    cols = {"Time": t, "Theta": rng.normal(size=T) * 0.1, "Gamma": rng.normal(size=T) * 0.1}
    d = rng.normal(size=(T, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    P1 = 1000 * d * rng.uniform(0.4, 3.1, T)[:, None]
    for j, ax in enumerate("XYZ"):
        cols[f"rod_end {ax}"] = rng.normal(size=T) * 5
        cols[f"robot_cable_attach_point {ax}"] = P1[:, j]
        cols[f"rob_cor_speed {ax}"] = rng.normal(size=T) * 80
    df = pd.DataFrame(cols)
    got = cb.build_theta_features(df, 3.0, 1.521)
    want = orc.build_theta_features(df, 3.0, 1.521)
    assert_close(got, want, 1e-9, 1e-12, "200k rows")


def test_resample_bit_exact(cb):
    pd = pytest.importorskip("pandas")
    g = golden("feature_layouts.npz")
    names = list(g["ref_resampled_columns"])
    df = pd.DataFrame({c: g[c] for c in names})
    for key, dt in (("ref_resampled", 0.05), ("ref_resampled_fine", 0.0125)):
        rs = cb.uniform_resample(df, dt)
        assert list(rs.columns) == names
        assert np.array_equal(rs.values, g[key]), key
    # extrapolation on both sides and exact knots
    t = np.array([0.0, 1.0, 2.5, 4.0])
    y = np.array([[1.0, -2.0], [3.0, 0.5], [2.0, 0.25], [8.0, 1.0]])
    tn = np.array([-1.0, 0.0, 1.0, 2.5, 4.0, 5.5, 3.3])
    assert np.array_equal(cb.resample_linear(t, y, tn), orc.resample_linear(t, y, tn))
    with pytest.raises(ValueError):
        cb.resample_linear(t[:1], y[:1], tn)
    with pytest.raises(ValueError):
        cb.uniform_resample(df.drop(columns=["Time"]))


def test_trajectory_csv_and_expression_in_rk4(cb, tmp_path):
    p = tmp_path / "rov_trajectory_exp1.csv"
    hdr = ", ".join(f"br{r}_{n}" for r in (0, 1) for n in "x y z phi theta psi u v w p q r".split())
    rows = np.arange(48, dtype=float).reshape(2, 24) / 8
    p.write_text(hdr + "\n" + "\n".join(",".join(f"{v:.3f}" for v in r) for r in rows) + "\n")
    tr = cb.load_trajectory_csv(str(p))
    assert tr.shape == (2, 2, 12) and tr[1, 0, 0] == rows[0, 12] and tr[0, 1, 11] == rows[1, 11]
    # rk4_integration accepts a run-time ExpressionModel in place of a generated EquationModel
    rng = np.random.default_rng(1)
    X = rng.normal(size=(300, 18))
    time = np.arange(300) * 0.05
    mdl = orc.load_model()
    row = next(r for r in mdl["dtheta_dt"]["rows"] if r["complexity"] == mdl["dtheta_dt"]["shipped_complexity"])
    a = cb.rk4_integration(cb.EquationModel("dtheta_dt"), X, time, 0.1)
    b = cb.rk4_integration(cb.ExpressionModel(row["sympy_format"], 18), X, time, 0.1)
    assert_close(b, a, 1e-12, 1e-15, "ExpressionModel through rk4_integration")


def test_config1_full_pipeline(cb):
    """T = 5,000 synthetic trajectory through features -> scaler -> RK4 of both ODEs ->
    per-step catenary parameter -> theta/gamma-augmented curve, against the oracle."""
    import importlib
    feat = importlib.import_module(cb.__name__ + ".features")
    cols = feat.config1_trajectory(5000)
    got = feat.config1_pipeline(cols)
    X = orc.scale_features(orc.extract_features(cols))
    t = cols["Time"]
    th = orc.rk4_integration_vec(orc.EquationModel("dtheta_dt"), X, t, cols["Theta"][0])
    ga = orc.rk4_integration_vec(orc.EquationModel("dgamma_dt"), X, t, cols["Gamma"][0])
    assert_close(got["theta"], th, 1e-9, 1e-12, "theta(t)")
    assert_close(got["gamma"], ga, 1e-9, 1e-12, "gamma(t)")
    P1 = np.stack([cols[f"robot_cable_attach_point {a}"] for a in "XYZ"], 1) / 1000
    l = np.hypot(P1[:, 0], P1[:, 1])
    assert_close(got["C"], orc.solve_catenary(l, P1[:, 2], 3.0), 1e-9, 0, "C(t)")
    assert got["curves"].shape == (5000, 100, 3) and got["valid"].all()
    cat = orc.Catenary(3.0)
    n_fallback = 0
    for i in range(0, 5000, 101):
        want = orc.transform_catenary(np.zeros(3), P1[i], cat, th[i], ga[i])[3]
        if want.shape[0] == 2:      # rotated span outside the solve domain: the reference falls back to [A, B'] (main_fun.py:67-69)
            assert not (got["valid"][i] & 2)
            assert_close(got["curves"][i][[0, -1]], want, 1e-9, 1e-12, f"chord fallback at step {i}")
            n_fallback += 1
            continue
        assert got["valid"][i] & 2
        assert_close(got["curves"][i], want, 1e-9, 1e-12, f"final curve at step {i}")
    assert n_fallback < 25
    # the reference's own per-step loop on a prefix (one predict per stage, simulate_rk4_theta_gamma.py:52-68)
    assert_close(got["theta"][:60], orc.rk4_integration(orc.EquationModel("dtheta_dt"), X[:60], t[:60], cols["Theta"][0]),
                 1e-9, 1e-12, "reference loop prefix")
