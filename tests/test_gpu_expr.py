"""Run-time equation evaluator on the GPU: every equation of the reference's PySR outputs
against the golden corpus (sympy.lambdify / Python-eval values), the integrators against the oracle."""
import importlib

import numpy as np
import pytest

from conftest import assert_close
from oracle import catenary_oracle as orc
from test_expr_cpu import close, corpus, corpus_inputs

pytestmark = pytest.mark.gpu


def test_every_reference_equation_on_gpu(cb):
    ex = importlib.import_module(cb.__name__ + ".expr")
    meta, y = corpus()
    bad = []
    for i, it in enumerate(meta["items"]):
        m = ex.ExpressionModel(it["expr"], it["n_features"], it["feature_names"])
        got = m.predict(corpus_inputs(meta, i))
        if not close(got, y[i], 1e-9):
            bad.append((i, it["expr"], got, y[i]))
    assert not bad, f"{len(bad)} of {len(meta['items'])} differ, first: {bad[0]}"


def test_large_batch_scaler_and_fp32(cb):
    ex = importlib.import_module(cb.__name__ + ".expr")
    mdl = orc.load_model()
    mean, scale = np.array(mdl["scaler_mean"]), np.array(mdl["scaler_scale"])
    s = "x3*(-0.05) + sin(x17)*0.048 - tanh(x0*x1)/(Abs(x2) + 1.5) + square(x16) - exp(-x5**2)"
    rng = np.random.default_rng(3)
    X = mean + rng.normal(size=(200_001, 18)) * scale
    want = orc.LambdifiedModel(s, 18, scaler_mean=mean, scaler_scale=scale).predict(X)
    m = ex.ExpressionModel(s, 18, scaler_mean=mean, scaler_scale=scale)
    assert_close(m.predict(X), want, 1e-9, 1e-12, "scaled batch")
    got32 = m.predict(X, dtype=np.float32)
    assert got32.dtype == np.float32
    assert np.abs(got32 - want).max() < 2e-4 * max(1.0, np.abs(want).max())
    # same answer as the build-time generated code for the shipped equation
    row = next(r for r in mdl["dtheta_dt"]["rows"] if r["complexity"] == mdl["dtheta_dt"]["shipped_complexity"])
    Xs = rng.normal(size=(5000, 18))
    gen = cb.EquationModel("dtheta_dt").predict(Xs)
    vm = ex.ExpressionModel(row["sympy_format"], 18).predict(Xs)
    assert_close(vm, gen, 1e-12, 1e-15, "VM vs generated code")


@pytest.mark.parametrize("method", ["rk4", "euler"])
def test_integrate_matches_oracle(cb, method):
    ex = importlib.import_module(cb.__name__ + ".expr")
    rng = np.random.default_rng(8)
    T, F = 5000, 11                                          # an 11-feature run (C6_2_1KIter_11f layout)
    s = "0.03*x4 - 0.01*sin(x9*2.5) + tanh(x1)*x10*0.02 - 0.004"
    X = rng.normal(size=(T, F))
    time = np.cumsum(rng.uniform(0.03, 0.07, T))             # non-uniform dt
    om = orc.LambdifiedModel(s, F)
    m = ex.ExpressionModel(s, F)
    if method == "rk4":
        want = orc.rk4_integration_vec(om, X, time, 0.3)
    else:
        want = orc.integrate_theta_gamma(om, om, X, time, 0.3, 0.3)[0]
    assert_close(m.integrate(X, time, 0.3, method=method), want, 1e-9, 1e-12, method)
    # batch of trajectories = the single-trajectory answers
    Xb = rng.normal(size=(7, 300, F))
    yb = m.integrate(Xb, time[:300], np.linspace(-0.2, 0.2, 7), method=method)
    for j in (0, 6):
        assert np.array_equal(yb[j], m.integrate(Xb[j], time[:300], np.linspace(-0.2, 0.2, 7)[j], method=method))
    # the reference's own loop (one predict per stage) on a short prefix
    if method == "rk4":
        assert_close(m.integrate(X[:40], time[:40], 0.3), orc.rk4_integration(om, X[:40], time[:40], 0.3), 1e-9,
                     1e-12, "reference loop")


def test_second_order_matches_oracle(cb):
    ex = importlib.import_module(cb.__name__ + ".expr")
    rng = np.random.default_rng(9)
    T = 3000
    names = ["theta", "dtheta", "v_surge", "a_x", "T_l"]    # named features as in the dd_* runs
    s = "-2.1*theta - 0.4*dtheta + 0.02*v_surge*cos(theta) + square(a_x)*0.001 - T_l/50"
    X = rng.normal(size=(T, 5))
    time = np.arange(T) * 0.05
    om = orc.LambdifiedModel(s, 5, names)
    m = ex.ExpressionModel(s, feature_names=names)
    want_y, want_rate = orc.integrate_second_order(om.predict(X), time, 0.12)
    y, rate = m.integrate(X, time, 0.12, method="second_order", return_rate=True)
    assert_close(rate, want_rate, 1e-9, 1e-12, "rate")
    assert_close(y, want_y, 1e-9, 1e-12, "angle")
    assert np.array_equal(ex.integrate_second_order(m, X, time, 0.12), y)


def test_expr_errors_and_edges(cb):
    ex = importlib.import_module(cb.__name__ + ".expr")
    m = ex.ExpressionModel("log(x0) + sqrt(x1) + x2/x3 + sign(x0)", 4)
    X = np.array([[-1.0, -4.0, 1.0, 0.0], [0.0, 4.0, 0.0, 0.0], [np.e, 9.0, -1.0, 0.0]])
    with np.errstate(all="ignore"):
        want = np.log(X[:, 0]) + np.sqrt(X[:, 1]) + X[:, 2] / X[:, 3] + np.sign(X[:, 0])
    got = m.predict(X)
    assert np.array_equal(np.isnan(got), np.isnan(want)) and np.array_equal(got[~np.isnan(want)], want[~np.isnan(want)])
    assert m.predict(np.zeros((0, 4))).shape == (0,)
    with pytest.raises(ValueError):
        m.predict(np.zeros((3, 5)))
    assert ex.ExpressionModel("2.5", 3).predict(np.zeros((4, 3))).tolist() == [2.5] * 4
    # a hand-made malformed program is an error value from the C ABI, not a crash
    bad = ex.Program([(2, 0)], [], 2)
    with pytest.raises(ValueError):
        ex.ExpressionModel(bad).predict(np.zeros((2, 2)))
    y = m.integrate(np.ones((1, 4)), np.zeros(1), 0.5)
    assert y.tolist() == [0.5]


def test_mpc_step_with_runtime_equations(cb):
    """The sampling-MPC step driven by equation STRINGS (a retrained model) instead of generated code."""
    rng = np.random.default_rng(21)
    K, H = 3000, 30
    sig = np.array(orc.load_model()["scaler_scale"][3:6])
    U = np.clip(rng.normal(size=(K, H, 3)), -3, 3) * sig
    pref = cb.circle_reference(H, 0.01)
    mdl = orc.load_model()
    ship_t = next(r for r in mdl["dtheta_dt"]["rows"] if r["complexity"] == mdl["dtheta_dt"]["shipped_complexity"])
    ship_g = next(r for r in mdl["dgamma_dt"]["rows"] if r["complexity"] == mdl["dgamma_dt"]["shipped_complexity"])
    # 1. the shipped pair as strings == the generated code path
    _, _, J_gen = cb.mpc_sample_step(cb.default_mpc_params(), U, pref)
    bi, bc, J_vm = cb.mpc_sample_step(cb.default_mpc_params(theta_expr=ship_t["sympy_format"],
                                                            gamma_expr=ship_g["sympy_format"]), U, pref)
    assert_close(J_vm, J_gen, 1e-11, 0, "VM vs generated (shipped pair)")
    assert bi == int(np.argmin(J_gen))
    # 2. equations that read the angles themselves (no k1 carry), other operators, against the oracle
    te = "-0.03*x16 - 0.8*x14 + 0.02*sin(x17) - 0.01*tanh(x3*0.5) + 0.004*square(x9)"
    ge = "0.5*(x15 - x17) - 0.3*x15*Abs(x13) + 0.002*cos(x4)"
    po = orc.default_mpc_params(theta_expr=te, gamma_expr=ge)
    Jo = orc.mpc_rollout_costs(po, U, pref)
    bi, bc, J = cb.mpc_sample_step(cb.default_mpc_params(theta_expr=te, gamma_expr=ge), U, pref)
    assert_close(J, Jo, 1e-9, 0, "VM rollout vs oracle")
    assert bi == int(np.argmin(Jo))
    # 3. ExpressionModel objects, SoA layout, fp32
    ex = importlib.import_module(cb.__name__ + ".expr")
    p2 = cb.default_mpc_params(theta_expr=ex.ExpressionModel(te, 18), gamma_expr=ex.ExpressionModel(ge, 18))
    b2, _, J2 = cb.mpc_sample_step(p2, np.ascontiguousarray(U.transpose(1, 2, 0)), pref, layout="H3K")
    assert b2 == bi and np.array_equal(J2, J)
    _, _, J32 = cb.mpc_sample_step(p2, U.astype(np.float32), pref, dtype=np.float32)
    assert_close(J32, Jo, 1e-4, 0, "VM rollout fp32")
    with pytest.raises(ValueError):
        cb.mpc_sample_step(cb.default_mpc_params(theta_expr=te), U, pref)
    with pytest.raises(ValueError):
        cb.mpc_sample_step(cb.default_mpc_params(theta_expr="x20", gamma_expr=ge), U, pref)
