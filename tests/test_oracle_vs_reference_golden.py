"""Pin the oracle: its restatements must reproduce the outputs of the reference's OWN functions stored in
tests/golden/*.npz (written by oracle/make_golden.py, which executes /root/reference code)."""
import numpy as np
import pytest

from conftest import assert_close, golden
from oracle import catenary_oracle as orc
from oracle import reference_loader as ref


def test_rodrigues_matches_reference():
    g = golden("rodrigues.npz")
    out = orc.rodrigues_rotation(g["v"], g["axis"], g["angle"])
    assert_close(out, g["ref_out"], 1e-13, 1e-15, "rodrigues")


def test_solve_catenary_matches_brentq_golden():
    g = golden("solve_catenary.npz")
    # brentq stops at xtol=2e-12, rtol=8.9e-16 -> the reference itself is only good to ~1e-11 relative
    assert_close(orc.solve_catenary(g["l"], g["dH"], 3.0), g["ref_C_L3"], 2e-11, 4e-12, "C (L=3)")
    assert_close(orc.solve_catenary(g["l"], g["dH"], g["L"]), g["ref_C"], 2e-11, 4e-12, "C (L varied)")
    # survey-time known answers from the unmodified reference
    kat = orc.solve_catenary(np.array([1.0, np.sqrt(2.0), 2.0, 2.9]), np.array([0.0, -1.0, 0.5, 0.0]), 3.0)
    assert_close(kat, [5.67689276, 3.07919405, 1.5916068, 0.3120942], 2e-8, 0, "KAT")
    assert np.isnan(g["ref_C_L3"]).sum() > 50        # the nan domain is exercised


def test_solve_catenary_exact_loop_matches_golden():
    g = golden("solve_catenary.npz")
    sl = slice(0, 80)
    assert_close(orc.solve_catenary_brentq(g["l"][sl], g["dH"][sl], 3.0), g["ref_C_L3"][sl], 1e-15, 0, "brentq loop")


def test_transform_catenary_matches_reference():
    g = golden("transform_catenary.npz")
    for i in range(g["A"].shape[0]):
        res = orc.transform_catenary(g["A"][i], g["B"][i], orc.Catenary(float(g["L"][i])), g["theta"][i],
                                     g["gamma"][i])
        for j in range(4):
            n = g["n_pts"][i][0 if j == 0 else 1]
            assert res[j].shape == (n, 3)
            assert_close(res[j], g["ref_curves"][i, j, :n], 1e-12, 1e-13, f"scenario {i} output {j}")
    assert (g["n_pts"] == 2).any() and (g["n_pts"] == 100).any()      # both the curve and the [A,B] fallback


def test_features_scaler_match_reference():
    g = golden("dynamics.npz")
    X = orc.extract_features(g)
    assert_close(X, g["ref_X"], 1e-13, 1e-13, "extract_features")
    assert_close(orc.scale_features(X), g["ref_Xs"], 1e-12, 1e-12, "StandardScaler.transform")


@pytest.mark.parametrize("nm", ["dtheta_dt", "dgamma_dt"])
def test_integrators_match_reference(nm):
    g = golden("dynamics.npz")
    m = orc.EquationModel(nm)
    assert_close(m.predict(g["ref_Xs"]), g[f"ref_predict_{nm}"], 0, 0, "predict")
    assert_close(orc.rk4_integration(m, g["ref_Xs"], g["Time"], 0.05), g[f"ref_rk4_{nm}"], 0, 0, "rk4 loop")
    assert_close(orc.rk4_integration_vec(m, g["ref_Xs"], g["Time"], 0.05), g[f"ref_rk4_{nm}"], 1e-13, 1e-15, "rk4 vec")
    assert_close(orc.rk4_integration_vec(m, g["ref_Xs"], g["time_nu"], -0.02), g[f"ref_rk4_nu_{nm}"], 1e-13, 1e-15,
                 "rk4 non-uniform")


def test_euler_matches_reference():
    g = golden("dynamics.npz")
    th, ga = orc.integrate_theta_gamma(orc.EquationModel("dtheta_dt"), orc.EquationModel("dgamma_dt"), g["ref_Xs"],
                                       g["Time"], 0.05, 0.1)
    assert_close(th, g["ref_euler_theta"], 1e-14, 1e-16, "euler theta")
    assert_close(ga, g["ref_euler_gamma"], 1e-14, 1e-16, "euler gamma")


def test_catenary_3d_matches_reference():
    g = golden("catenary_3d.npz")
    for i in range(4):
        assert_close(orc.compute_catenary_3D(g["p0"][i], g["p1"][i], float(g["L"][i]), 100), g["ref_pts"][i], 1e-13,
                     1e-13, f"catenary_3d {i}")


def test_standin_catenary_physics():
    """The stand-in for the missing pympc class (parity unpinned upstream) must at least be a catenary:
    end points exact, arc length == cable length, sag toward -z."""
    cat = orc.Catenary(3.0, "ENU")
    a, b = np.zeros(3), np.array([1.0, 1.0, -1.0])
    C, H, D, pts = cat(a, b)
    fine = orc.Catenary(3.0, "ENU", n_points=20001)(a, b)[3]
    arc = np.linalg.norm(np.diff(fine, axis=0), axis=1).sum()
    assert abs(arc - 3.0) < 1e-6
    assert np.allclose(pts[0], a) and np.allclose(pts[-1], b, atol=1e-12)
    assert pts[:, 2].min() < -1.0 and H > 0 and 0 <= D <= np.hypot(1, 1)
    assert abs((b[2] - H) - fine[:, 2].min()) < 1e-6
    assert cat(a, np.array([2.5, 2.0, 0.5]))[3] is None          # taut -> None (main_fun.py:67-69 fallback)


@pytest.mark.skipif(not ref.available(), reason="/root/reference is only present in the build container")
def test_live_reference_spot_check():
    mf = ref.main_fun()
    rng = np.random.default_rng(5)
    l, dH = rng.uniform(0.3, 3.0, 40), rng.uniform(-2, 2, 40)
    with np.errstate(all="ignore"):
        assert_close(orc.solve_catenary(l, dH, 3.0), mf.solve_catenary(l, dH, 3.0), 2e-11, 4e-12, "live solve")
    v, k = rng.normal(size=3), rng.normal(size=3)
    assert_close(orc.rodrigues_rotation(v, k, 0.7), mf.rodrigues_rotation(v, k, 0.7), 1e-14, 1e-16, "live rodrigues")


def test_philox_known_answer():
    """Philox4x32-10 KAT from Random123 (counter = key = 0 and the all-ones vector)."""
    r = orc.philox4x32_10(0, 0, 0, 0, 0, 0)
    assert [int(x) for x in r] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    r = orc.philox4x32_10(0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff)
    assert [int(x) for x in r] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
