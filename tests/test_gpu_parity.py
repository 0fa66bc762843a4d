"""Parity of the CUDA path (through the C ABI) against the oracle and the reference-generated golden vectors.
Tolerances: fp64 1e-9 relative (north_star), fp32 fast mode 1e-4, argmin index exact in fp64."""
import numpy as np
import pytest

from conftest import assert_close, golden
from oracle import catenary_oracle as orc

pytestmark = pytest.mark.gpu
RTOL64 = 1e-9
RTOL32 = 1e-4
SIGMA = np.array(orc.load_model()["scaler_scale"][3:6])      # config 3: sigma of V1 (mm/s)


# ---------------------------------------------------------------- geometry
def test_solve_catenary_golden(cb):
    g = golden("solve_catenary.npz")
    assert_close(cb.solve_catenary(g["l"], g["dH"], 3.0), g["ref_C_L3"], 2e-11, 4e-12, "C vs reference brentq")
    assert_close(cb.solve_catenary(g["l"], g["dH"], g["L"]), g["ref_C"], 2e-11, 4e-12, "C, L per element")
    c = cb.solve_catenary(1.0, 0.0, 3.0)
    assert c.shape == () and abs(float(c) - 5.67689276) < 1e-7            # scalar call, np.vectorize shape


def test_solve_catenary_oracle_large_and_edges(cb):
    rng = np.random.default_rng(0)
    n = 200_000
    l = rng.uniform(1e-3, 3.3, n)
    dH = rng.uniform(-3.1, 3.1, n)
    assert_close(cb.solve_catenary(l, dH, 3.0), orc.solve_catenary(l, dH, 3.0), RTOL64, 0, "C 200k")
    # empty, zero span, taut, |dH| >= L
    assert cb.solve_catenary(np.zeros(0), np.zeros(0), 3.0).shape == (0,)
    edge = cb.solve_catenary(np.array([0.0, 3.0, 1.0, 1.0, 2.999999]), np.array([0.0, 0.0, 3.0, -3.5, 0.0]), 3.0)
    assert np.isnan(edge[:4]).all() and np.isfinite(edge[4])
    with np.errstate(all="ignore"):
        assert_close(edge, orc.solve_catenary(np.array([0.0, 3.0, 1.0, 1.0, 2.999999]),
                                              np.array([0.0, 0.0, 3.0, -3.5, 0.0]), 3.0), 1e-7, 0, "edges")


def test_solve_catenary_fp32(cb):
    """fp32 fast mode over the reference's whole domain (main_fun.py:418-431), spans arbitrarily close to taut
    included: 1e-4 relative against the oracle on the SAME float32-rounded inputs.  (N = L^2 - dH^2 - l^2 is formed in
    fp64 inside the float kernel; in pure fp32 this was 3e-4 within 0.1% of taut.)"""
    rng = np.random.default_rng(1)
    n = 400_000
    l = rng.uniform(1e-3, 3.3, n).astype(np.float32)
    dH = rng.uniform(-3.1, 3.1, n).astype(np.float32)
    # a sub-population within 1e-3 ... 1e-6 (relative) of taut: l = sqrt(L^2 - dH^2) (1 - eps)
    m = n // 10
    eps = 10.0 ** rng.uniform(-6, -3, m)
    l[:m] = (np.sqrt(9.0 - dH[:m].astype(np.float64) ** 2) * (1 - eps)).astype(np.float32)
    got = cb.solve_catenary(l, dH, 3.0, dtype=np.float32).astype(np.float64)
    want = orc.solve_catenary(l.astype(np.float64), dH.astype(np.float64), 3.0)
    # the nan domain may differ only where C sits on a bracket end (1e-6 or 10) to within the fp32 tolerance
    interior = ~np.isnan(want) & (want > 1e-6 * (1 + 1e-3)) & (want < 10.0 * (1 - 1e-3))
    assert interior.sum() > 150_000 and not np.isnan(got[interior]).any()
    rel = np.abs(got[interior] / want[interior] - 1)
    assert rel.max() < RTOL32, f"fp32 C worst relative error {rel.max():.3e}"
    taut = interior[:m]
    assert taut.sum() > 1000 and np.abs(got[:m][taut] / want[:m][taut] - 1).max() < RTOL32
    outside = np.isnan(want) & ~np.isnan(got)
    assert outside.mean() < 1e-4                              # bracket-end flips only


def test_solve_catenary_residual_property(cb):
    """Size-independent property at config-2 scale: the returned root satisfies the reference's equation."""
    rng = np.random.default_rng(2)
    n = 1 << 20
    l = rng.uniform(0.4, 2.9, n)
    dH = rng.uniform(-1.5, 1.5, n)
    C = cb.solve_catenary(l, dH, 3.0)
    ok = ~np.isnan(C)
    assert ok.mean() > 0.5
    arc = 2 * np.sinh(0.5 * l[ok] * C[ok]) / C[ok]
    assert np.abs(arc / np.sqrt(9.0 - dH[ok] ** 2) - 1).max() < 1e-12


def test_rodrigues_golden(cb):
    g = golden("rodrigues.npz")
    assert_close(cb.rodrigues_rotation(g["v"], g["axis"], g["angle"]), g["ref_out"], 1e-12, 1e-14, "rodrigues batch")
    one = cb.rodrigues_rotation(g["v"][3], g["axis"][3], g["angle"][3])
    assert one.shape == (3,)
    assert_close(one, g["ref_out"][3], 1e-12, 1e-14, "rodrigues single")


def test_transform_catenary_golden(cb):
    g = golden("transform_catenary.npz")
    for i in range(g["A"].shape[0]):
        res = cb.transform_catenary(g["A"][i], g["B"][i], cb.Catenary(float(g["L"][i])), g["theta"][i], g["gamma"][i])
        for j in range(4):
            n = g["n_pts"][i][0 if j == 0 else 1]
            assert res[j].shape == (n, 3), (i, j)
            assert_close(res[j], g["ref_curves"][i, j, :n], RTOL64, 1e-12, f"scenario {i} output {j}")


def test_transform_catenary_foreign_callable(cb):
    """Any catenary_fn works (main_fun.py:64-69): curves from the callable, rotations on the device."""
    g = golden("transform_catenary.npz")
    for i in (1, 5, 6):
        res = cb.transform_catenary(g["A"][i], g["B"][i], orc.Catenary(float(g["L"][i])), g["theta"][i], g["gamma"][i])
        for j in range(4):
            n = g["n_pts"][i][0 if j == 0 else 1]
            assert_close(res[j], g["ref_curves"][i, j, :n], RTOL64, 1e-12, f"foreign fn {i}/{j}")


def test_catenary_class_surface(cb):
    a, b = np.zeros(3), np.array([1.0, 1.0, 0.0])          # catenary.py:13-14
    out = cb.Catenary(length=3.0, reference_frame="ENU")(a, b)
    ref = orc.Catenary(3.0, "ENU")(a, b)
    assert len(out) == 4 and out[3].shape == (100, 3)
    assert_close(out[3], ref[3], RTOL64, 1e-13, "points")
    assert_close(out[:3], ref[:3], RTOL64, 1e-12, "C,H,D")
    assert cb.Catenary(3.0)(a, np.array([2.5, 2.0, 0.5]))[3] is None
    ned = cb.Catenary(3.0, "NED")(a, np.array([1.0, 1.0, 1.0]))
    assert_close(ned[3], orc.Catenary(3.0, "NED")(a, np.array([1.0, 1.0, 1.0]))[3], RTOL64, 1e-13, "NED")
    idx, pt = cb.lowest_point(out[3])                       # fully_augmented_catenary.py:20-22
    assert idx == int(np.argmin(out[3][:, 2])) and np.array_equal(pt, out[3][idx])   # np.argmin of the SAME curve
    # level span: samples 49 and 50 tie in exact arithmetic, so only the depth is comparable with the oracle
    assert abs(pt[2] - ref[3][:, 2].min()) < 1e-12
    b2 = np.array([1.0, 1.0, -1.0])                         # fully_augmented_catenary.py:12: unique lowest sample
    c2, r2 = cb.Catenary(3.0)(a, b2)[3], orc.Catenary(3.0)(a, b2)[3]
    assert cb.lowest_point(c2)[0] == int(np.argmin(r2[:, 2]))
    with pytest.raises(ValueError):
        cb.Catenary(3.0, "XYZ")


def test_shapes_batch_vs_oracle(cb):
    rng = np.random.default_rng(0)
    n = 4000
    A = rng.normal(size=(n, 3)) * 0.1
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    dist = rng.uniform(0.3, 2.85, n)
    L = rng.uniform(2.0, 5.0, n)
    dist[:40] = L[:40] * rng.uniform(1.0, 1.05, 40)         # taut -> invalid
    d[40:80, :2] *= 1e-8                                     # near vertical
    B = A + d * dist[:, None]
    th, ga = rng.uniform(-0.52, 0.52, n), rng.uniform(-0.52, 0.52, n)
    curves, C, valid = cb.catenary_shapes(A, B, L, th, ga, which=15)
    rel = B - A
    assert_close(C, orc.solve_catenary(np.hypot(rel[:, 0], rel[:, 1]), rel[:, 2], L), RTOL64, 0, "C")
    assert (valid[:40] & 1 == 0).all()
    for i in range(0, n, 29):
        ref = orc.transform_catenary(A[i], B[i], orc.Catenary(L[i]), th[i], ga[i])
        for j in range(4):
            got = curves[j][i] if ref[j].shape[0] == 100 else curves[j][i][[0, -1]]
            assert_close(got, ref[j], RTOL64, 1e-12, f"cfg {i} out {j}")
    # only 'final' requested == the 4th block of the full request; fp32 mode within 1e-4
    (fin,), _, _ = cb.catenary_shapes(A, B, L, th, ga)
    assert np.array_equal(fin, curves[3])
    (f32,), _, v32 = cb.catenary_shapes(A[80:], B[80:], L[80:], th[80:], ga[80:], dtype=np.float32)
    good = (valid[80:] == 3) & (v32 == 3)
    assert good.mean() > 0.5
    assert np.abs(f32[good] - fin[80:][good]).max() < 2e-4 * 5


def test_shapes_config2_scale_properties(cb):
    """Config 2 (1M x 100) through the chunked host pipeline: endpoints pinned, arc length == L."""
    rng = np.random.default_rng(0)
    n = 1 << 20
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    B = d * rng.uniform(1.2, 2.85, n)[:, None]
    A = np.zeros((n, 3))
    (orig,), C, valid = cb.catenary_shapes(A, B, 3.0, which=1)
    ok = (valid & 1) == 1
    assert ok.mean() > 0.5
    assert np.abs(orig[:, 0]).max() == 0 and np.abs(orig[:, -1] - B).max() < 1e-12
    sel = np.flatnonzero(ok)[::1000]
    seg = np.linalg.norm(np.diff(orig[sel], axis=1), axis=2).sum(1)
    assert (seg < 3.0 + 1e-9).all() and (seg > 2.97).all()      # 100-point polyline slightly shorter than the cable
    (fin,), _, _ = cb.catenary_shapes(A[sel], B[sel], 3.0, np.zeros(sel.size), np.zeros(sel.size))
    assert np.abs(fin - orig[sel]).max() < 1e-12                # theta = gamma = 0 is the identity


# ---------------------------------------------------------------- dynamics
def test_features_golden(cb):
    g = golden("dynamics.npz")
    assert_close(cb.extract_features(g), g["ref_X"], RTOL64, 1e-12, "extract_features")
    assert_close(cb.extract_features(g, scaled=True), g["ref_Xs"], RTOL64, 1e-10, "scaled features")


@pytest.mark.parametrize("nm", ["dtheta_dt", "dgamma_dt"])
def test_integrators_golden(cb, nm):
    g = golden("dynamics.npz")
    m = cb.EquationModel(nm)
    assert_close(m.predict(g["ref_Xs"]), g[f"ref_predict_{nm}"], RTOL64, 1e-13, "predict")
    assert_close(cb.rk4_integration(m, g["ref_Xs"], g["Time"], 0.05), g[f"ref_rk4_{nm}"], RTOL64, 1e-12, "rk4")
    assert_close(cb.rk4_integration(m, g["ref_Xs"], g["time_nu"], -0.02), g[f"ref_rk4_nu_{nm}"], RTOL64, 1e-12,
                 "rk4 non-uniform dt")
    y32 = cb.rk4_integration(m, g["ref_Xs"], g["Time"], 0.05, dtype=np.float32)
    assert np.abs(y32 - g[f"ref_rk4_{nm}"]).max() < 1e-4


def test_euler_golden_and_edges(cb):
    g = golden("dynamics.npz")
    mt, mg = cb.EquationModel("dtheta_dt"), cb.EquationModel("dgamma_dt")
    th, ga = cb.integrate_theta_gamma(mt, mg, g["ref_Xs"], g["Time"], 0.05, 0.1)
    assert_close(th, g["ref_euler_theta"], RTOL64, 1e-12, "euler theta")
    assert_close(ga, g["ref_euler_gamma"], RTOL64, 1e-12, "euler gamma")
    one = cb.rk4_integration(mt, g["ref_Xs"][:1], g["Time"][:1], 0.3)          # T = 1: just y0
    assert one.shape == (1,) and one[0] == 0.3
    with pytest.raises(ValueError):
        cb.rk4_integration(mt, g["ref_Xs"][:, :17], g["Time"], 0.0)


def test_config1_rollout_and_batch(cb):
    """Config 1: 5,000-step open-loop RK4 of both ODEs, and 64 trajectories at once."""
    rng = np.random.default_rng(42)
    X = rng.normal(size=(64, 5000, 18))
    t = np.arange(5000) * 0.05
    for nm in ("dtheta_dt", "dgamma_dt"):
        y = cb.rk4_integration(cb.EquationModel(nm), X, t, np.linspace(-0.1, 0.1, 64))
        for i in (0, 63):
            want = orc.rk4_integration_vec(orc.EquationModel(nm), X[i], t, np.linspace(-0.1, 0.1, 64)[i])
            assert_close(y[i], want, RTOL64, 1e-11, f"config1 {nm} traj {i}")


def test_every_hall_of_fame_equation(cb):
    rng = np.random.default_rng(3)
    X = rng.normal(size=(2000, 18))
    tab = cb.equation_table()
    assert len(tab) == 37
    for i, nm, cx, live in tab:
        assert_close(cb.predict(X, i), orc.equation(nm, cx)(X), RTOL64, 1e-13, f"{nm} c{cx}")
    raw = rng.normal(size=(100, 18)) * np.array(orc.load_model()["scaler_scale"]) + np.array(orc.load_model()["scaler_mean"])
    i = cb.equation_id("dtheta_dt")
    assert_close(cb.predict(raw, i, apply_scaler=True), orc.equation("dtheta_dt")(orc.scale_features(raw)), RTOL64,
                 1e-12, "apply_scaler")


# ---------------------------------------------------------------- sampling MPC
def _controls(K, H, seed=1234):
    rng = np.random.default_rng(seed)
    return np.clip(rng.normal(size=(K, H, 3)), -3, 3) * SIGMA


def test_mpc_small_all_layouts(cb):
    K, H = 3000, 50                                         # ragged: not a multiple of the block size
    U = _controls(K, H)
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    Jo, traj = orc.mpc_rollout_costs(orc.default_mpc_params(), U, pref, return_traj=True)
    bi, bc, J = cb.mpc_sample_step(p, U, pref)
    assert_close(J, Jo, RTOL64, 0, "costs AoS")
    assert bi == int(np.argmin(Jo)) and bc == J[bi]
    bi2, bc2, J2 = cb.mpc_sample_step(p, np.ascontiguousarray(U.transpose(1, 2, 0)), pref, layout="H3K")
    assert np.array_equal(J, J2) and bi2 == bi              # layout does not change a single bit
    bi3, _, none = cb.mpc_sample_step(p, U, pref, return_costs=False)
    assert none is None and bi3 == bi
    b32, c32, J32 = cb.mpc_sample_step(p, U.astype(np.float32), pref, dtype=np.float32)
    assert_close(J32, Jo, RTOL32, 0, "costs f32")


def test_mpc_config3_full_size(cb):
    """Config 3: 65,536 candidates x horizon 50 -- costs within 1e-9, argmin index exact."""
    K, H = 65536, 50
    U = _controls(K, H)
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    Jo = orc.mpc_rollout_costs(orc.default_mpc_params(), U, pref)
    bi, bc, J = cb.mpc_sample_step(p, U, pref)
    assert_close(J, Jo, RTOL64, 0, "config 3 costs")
    assert bi == int(np.argmin(Jo))
    assert abs(bc - Jo[bi]) <= RTOL64 * abs(Jo[bi])


def test_mpc_constraints_active(cb):
    """Poses that trigger every cost term: cable nearly taut (length penalty, nan C penalty) and a deep sag
    below the seabed clearance."""
    K, H = 2048, 30
    U = _controls(K, H, 7) * 4.0
    for over in (dict(P1=(1.9, 1.9, -0.9)), dict(P1=(0.4, 0.3, -0.6), z_floor=-1.2), dict(P1=(1.0, 1.0, -1.0), L=2.0),
                 dict(P1=(0.05, 0.0, -1.0)), dict(dt=0.05)):
        po = orc.default_mpc_params(**over)
        pref = cb.circle_reference(H, po["dt"])
        Jo = orc.mpc_rollout_costs(po, U, pref)
        bi, bc, J = cb.mpc_sample_step(cb.default_mpc_params(**over), U, pref)
        assert_close(J, Jo, RTOL64, 0, f"costs {over}")
        assert bi == int(np.argmin(Jo)), over


def test_mpc_sine_arguments_beyond_fast_range(cb):
    """Commanded speeds of ~1e7 mm/s put the scaled V1x far beyond the |x| < 1e5 range of the rollout's own
    sine, mixed lane by lane with ordinary candidates: the deferred full-range re-evaluation must reproduce
    NumPy's sin.  Only the angle term is weighted, so the cost is the integral of theta^2 + gamma^2 and every
    stage slope matters; gamma_prev != gamma makes gamma (hence x17) move as well."""
    K, H = 1024, 24
    rng = np.random.default_rng(77)
    U = _controls(K, H, 5)
    far = rng.random(K) < 0.3
    U[far, :, 0] = rng.uniform(-3e7, 3e7, size=(int(far.sum()), H))
    U[::97, 3, 0] = 1.0e5 * 108.48977412955092 + 80.85390753943355   # scaled x3 lands at the threshold itself
    over = dict(w_trk=0.0, w_du=0.0, w_len=0.0, w_sag=0.0, w_ang=1.0, nan_penalty=0.0,
                gamma_prev=float(np.radians(20.5)))
    po = orc.default_mpc_params(**over)
    pref = cb.circle_reference(H, po["dt"])
    Jo = orc.mpc_rollout_costs(po, U, pref)
    bi, bc, J = cb.mpc_sample_step(cb.default_mpc_params(**over), U, pref)
    assert np.isfinite(Jo).all()
    assert_close(J, Jo, RTOL64, 0, "costs with far sine arguments")
    assert bi == int(np.argmin(Jo))


def test_mpc_other_equations(cb):
    K, H = 1024, 40
    U = _controls(K, H, 3)
    pref = cb.circle_reference(H, 0.01)
    for tcx, gcx in ((19, 7), (8, 5), (25, 9)):
        try:
            po = orc.default_mpc_params(theta_cx=tcx, gamma_cx=gcx)
            Jo = orc.mpc_rollout_costs(po, U, pref)
        except StopIteration:
            continue
        _, _, J = cb.mpc_sample_step(cb.default_mpc_params(theta_cx=tcx, gamma_cx=gcx), U, pref)
        assert_close(J, Jo, RTOL64, 0, f"eq pair {tcx},{gcx}")


def test_mpc_argmin_ties_and_nan(cb):
    """Winner order: duplicates -> first index (identical candidates across blocks); NaN costs rank last."""
    K, H = 1000, 10
    U = _controls(K, H, 9)
    U[900] = U[17]
    U[500] = U[17]
    p = cb.default_mpc_params(w_trk=0.0, w_ang=0.0, w_len=0.0, w_sag=0.0)      # cost = control effort only
    pref = cb.circle_reference(H, 0.01)
    U[17] *= 0.0
    U[500] *= 0.0
    U[900] *= 0.0
    bi, bc, J = cb.mpc_sample_step(p, U, pref)
    assert J[17] == J[500] == J[900] == 0.0 and bi == 17 and bc == 0.0
    U[300, 3, 1] = np.nan
    bi, bc, J = cb.mpc_sample_step(p, U, pref)
    assert np.isnan(J[300]) and bi == 17 and bc == 0.0        # a NaN cost never wins (oracle.argmin_mpc) ...
    assert (bi, bc) == orc.argmin_mpc(J)
    Un = np.full((200, H, 3), np.nan)
    bi, bc, J = cb.mpc_sample_step(p, Un, pref)
    assert bi == 0 and np.isnan(bc) and np.isnan(J).all()     # ... and with every cost NaN the lowest index does


def test_mpc_device_rng_matches_oracle_philox(cb):
    K, H = 4096, 50
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    Uo = orc.sample_controls(1234, np.arange(K), H, SIGMA)
    Jo = orc.mpc_rollout_costs(orc.default_mpc_params(), Uo, pref)
    bi, bc, J = cb.mpc_sample_step(p, None, pref, K=K, sigma=SIGMA, seed=1234)
    assert_close(J, Jo, RTOL64, 0, "rng costs")
    assert bi == int(np.argmin(Jo))
    # shards of the same job (k_offset) reproduce slices of it: the multi-GPU path on one GPU
    winners = []
    for lo in range(0, K, 1024):
        i, c, Js = cb.mpc_sample_step(p, None, pref, K=1024, sigma=SIGMA, seed=1234, k_offset=lo)
        assert np.array_equal(Js, J[lo:lo + 1024])
        winners.append((c, i))
    gi, gc = cb.reduce_winners([w[0] for w in winners], [w[1] for w in winners])
    assert gi == bi and gc == bc


def test_device_sampler_values_match_oracle(cb):
    """The sampled controls themselves (not only the costs they lead to): Philox words, the integer-keyed
    -2 ln u and sin/cos 2 pi u of the fp64 sampler against NumPy's log/sin/cos, incl. high candidate indices
    (k >= 2^32) and a non-zero problem id.  1e-13 relative: the transcendental parts are good to ~3e-16."""
    import importlib
    dev = importlib.import_module(cb.__name__ + ".device")
    p = cb.default_mpc_params()
    for K, H, seed, k0, prob in ((8192, 50, 1234, 0, 0), (4096, 100, 2 ** 40 + 7, 2 ** 33 + 5, 9)):
        U = dev.sample_controls(p, K, H, SIGMA, seed=seed, k_offset=k0, problem_offset=prob)
        assert U.shape == (1, H, 3, K)
        Uo = orc.sample_controls(seed, k0 + np.arange(K, dtype=np.uint64), H, SIGMA, problem=prob)   # (K,H,3)
        assert_close(U[0].transpose(2, 0, 1), Uo, 1e-13, 1e-11, f"sampled controls seed {seed}")   # atol: 2 pi u is rounded in the oracle
        assert np.all(np.abs(U[0]) <= 3.0 * np.asarray(SIGMA)[None, :, None] * (1 + 1e-15))


def test_mpc_fleet(cb):
    """Config 5 semantics at reduced size: independent problems, segmented argmin, no cross-talk."""
    n, K, H = 6, 1500, 20
    rng = np.random.default_rng(10_000)
    p = cb.default_mpc_params()
    poses = np.array([1.0, 1.0, -1.0]) + rng.normal(size=(n, 3)) * 0.15
    prm = dict(p, P1=poses, theta=rng.uniform(-0.2, 0.2, n), gamma=rng.uniform(-0.2, 0.2, n))
    prm["theta_prev"], prm["gamma_prev"] = prm["theta"], prm["gamma"]
    problems = cb.make_problems(prm, n)
    pref = np.stack([cb.circle_reference(H, p["dt"], center=poses[i]) for i in range(n)])
    bi, bc, J = cb.mpc_fleet_step(p, problems, pref, K, SIGMA, seed=77, return_costs=True, problem_offset=3)
    for i in range(n):
        Uo = orc.sample_controls(77, np.arange(K), H, SIGMA, problem=3 + i)
        po = orc.default_mpc_params(P1=poses[i], theta=prm["theta"][i], gamma=prm["gamma"][i],
                                    theta_prev=prm["theta"][i], gamma_prev=prm["gamma"][i])
        Jo = orc.mpc_rollout_costs(po, Uo, pref[i])
        assert_close(J[i], Jo, RTOL64, 0, f"fleet problem {i}")
        assert bi[i] == int(np.argmin(Jo)) and bc[i] == J[i, bi[i]]


def test_mpc_config4_shard_property(cb):
    """Large-K property (config-4 shard size, H=100): the device argmin equals np.argmin of the device costs,
    and a sampled subset of candidates matches the oracle."""
    K, H = 524288, 100
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    bi, bc, J = cb.mpc_sample_step(p, None, pref, K=K, sigma=SIGMA, seed=5, k_offset=3 * K)
    assert bi - 3 * K == int(np.argmin(J)) and bc == J[bi - 3 * K]
    pick = np.arange(0, K, 997)
    Uo = orc.sample_controls(5, 3 * K + pick, H, SIGMA)
    assert_close(J[pick], orc.mpc_rollout_costs(orc.default_mpc_params(), Uo, pref), RTOL64, 0, "sampled candidates")


def test_device_step_matches_host_call(cb):
    import importlib
    dev = importlib.import_module(cb.__name__ + ".device")
    K, H = 8192, 50
    U = np.ascontiguousarray(_controls(K, H).transpose(1, 2, 0))
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    bi, bc, J = cb.mpc_sample_step(p, U, pref, layout="H3K")
    st = dev.MpcDeviceStep(p, pref, K, controls=U, keep_costs=True)
    for _ in range(3):                                      # relaunch: tickets reset by the kernel itself
        st.launch()
    i2, c2 = st.result()
    assert int(i2[0]) == bi and float(c2[0]) == bc
    assert np.array_equal(st.costs.cpu().numpy()[0], J)
    st.launch(to_host=True)                                 # winner written by the kernel into mapped pinned memory
    i3, c3 = st.result_host()
    assert int(i3[0]) == bi and float(c3[0]) == bc


def test_errors_are_values_not_crashes(cb):
    p = cb.default_mpc_params()
    with pytest.raises(ValueError):
        cb.mpc_sample_step(p, np.zeros((4, 5, 2)), cb.circle_reference(5, 0.01))
    with pytest.raises(ValueError):
        cb.mpc_sample_step(dict(p, dt=-1.0), np.zeros((4, 5, 3)), cb.circle_reference(5, 0.01))
    with pytest.raises(ValueError):
        cb.catenary_shapes(np.zeros((2, 3)), np.zeros((3, 3)), 3.0)
    assert cb.default_context().launch_count > 0


def test_mpc_host_step_chunked_pipeline(cb):
    """K >= 32,768 with explicit controls goes up in 8 chunks (copy/rollout overlap) and the chunk winners fold on
    the device: same costs and the same first-minimum as one launch, in both layouts, ragged K, ties and NaN
    placed in different chunks."""
    K, H = 40_037, 12
    U = _controls(K, H, 77)
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    U[39_000] = U[123]                                      # identical candidates in the last and the first chunk
    U[20_500] = U[123]
    Jo = orc.mpc_rollout_costs(orc.default_mpc_params(), U, pref)
    bi, bc, J = cb.mpc_sample_step(p, U, pref)                                          # AoS, chunked
    assert_close(J, Jo, RTOL64, 0, "chunked AoS costs")
    assert bi == int(np.argmin(Jo)) and bc == J[bi]
    Us = np.ascontiguousarray(U.transpose(1, 2, 0))
    bi2, bc2, J2 = cb.mpc_sample_step(p, Us, pref, layout="H3K")                        # SoA, chunked (2-D copies)
    assert bi2 == bi and bc2 == bc and np.array_equal(J2, J)
    # make the duplicated candidate the global winner: the first of the three copies must win
    U0 = np.zeros_like(U) + U[123]
    best = orc.mpc_rollout_costs(orc.default_mpc_params(), U0[:1], pref)[0]
    Ubig = U.copy()
    Ubig[:, :, 0] += 4000.0                                  # everybody else far worse (tracking + du terms)
    for k in (123, 20_500, 39_000):
        Ubig[k] = U[123]
    b3, c3, J3 = cb.mpc_sample_step(p, Ubig, pref)
    assert b3 == 123 and abs(c3 - best) <= RTOL64 * best
    b4, _, _ = cb.mpc_sample_step(p, np.ascontiguousarray(Ubig.transpose(1, 2, 0)), pref, layout="H3K")
    assert b4 == 123
    Ubig[30_001, 3, 1] = np.nan                              # a NaN cost in a later chunk never wins
    b5, c5, J5 = cb.mpc_sample_step(p, Ubig, pref)
    assert np.isnan(J5[30_001]) and b5 == 123 and c5 == c3
    # k_offset shifts the reported index of the chunked call too
    b6, _, _ = cb.mpc_sample_step(p, U, pref, k_offset=1_000_000, return_costs=False)
    assert b6 == bi + 1_000_000
    # fp32 storage + compute through the same pipeline
    b7, c7, J7 = cb.mpc_sample_step(p, U.astype(np.float32), pref, dtype=np.float32)
    assert_close(J7, Jo, RTOL32, 0, "chunked fp32 costs")


def test_mpc_argmin_near_ties(cb):
    """Index exactness margin: pairs of candidates whose ORACLE costs differ by 1e-8 ... 1e-15 relative, the better one
    at the higher index and in another block/chunk (oracle/near_ties.py).  The kernel's costs are not bit-identical to
    NumPy's, so the index is exact down to the rounding noise of the two 50-step sums; the test pins that margin at
    1e-12 and prints what happens below it."""
    from oracle import near_ties
    K, H = 4096, 50
    U = _controls(K, H)
    p = cb.default_mpc_params()
    po = orc.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    results, lines = [], []
    for h_perturb in (0, H - 1):                             # a perturbation that propagates through all 50 steps / none
        for case in near_ties.plant_near_ties(po, U, pref, h_perturb=h_perturb, seed=h_perturb):
            Jo = orc.mpc_rollout_costs(po, case["U"], pref)
            assert orc.argmin_mpc(Jo)[0] == case["want"]
            bi, bc, J = cb.mpc_sample_step(p, case["U"], pref)                               # AoS, one launch
            bi2, _, _ = cb.mpc_sample_step(p, np.ascontiguousarray(case["U"].transpose(1, 2, 0)), pref, layout="H3K",
                                           return_costs=False)
            ok = bi == case["want"] and bi2 == case["want"]
            results.append((case["gap"], ok))
            lines.append(f"h={h_perturb} level {case['level']:.0e} ({case['better']} better) gap {case['gap']:.3e}: "
                         f"{'exact' if ok else 'MISMATCH'}  max|J/Jo-1| {np.abs(J / Jo - 1).max():.2e}")
    print("\n".join(lines))
    m = near_ties.margin(results)
    print("near-tie margin:", m)
    assert all(ok for gap, ok in results if gap >= 0.9e-12), m


def _config2_batch(n, seed, sweep_L):
    """P0 = 0, direction uniform on the sphere, d ~ U[0.3, 2.85], 1% taut (d ~ U[3.0, 3.2]),
    1% near-vertical (l < 1e-6), theta, gamma ~ U[-30, 30] deg; second sweep L ~ U[2, 5]."""
    rng = np.random.default_rng(seed)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    dist = rng.uniform(0.3, 2.85, n)
    n1 = max(n // 100, 1)
    dist[:n1] = rng.uniform(3.0, 3.2, n1)
    d[n1:2 * n1, :2] *= 1e-7 / np.maximum(np.hypot(d[n1:2 * n1, 0], d[n1:2 * n1, 1]), 1e-3)[:, None] * \
        rng.uniform(1e-3, 1.0, n1)[:, None]
    d[n1:2 * n1] /= np.linalg.norm(d[n1:2 * n1], axis=1, keepdims=True)
    A = np.zeros((n, 3))
    B = d * dist[:, None]
    L = rng.uniform(2.0, 5.0, n) if sweep_L else np.full(n, 3.0)
    th = np.radians(rng.uniform(-30, 30, n))
    ga = np.radians(rng.uniform(-30, 30, n))
    return A, B, L, th, ga


@pytest.mark.parametrize("sweep_L", [False, True])
def test_shapes_config2_distribution_all_curves(cb, sweep_L):
    """EVERY curve of 16,384 configurations drawn like BASELINE config 2 (taut and near-vertical sub-populations
    included), all four transform_catenary outputs: fp64 at 1e-9 against the batched oracle, and the fp32 fast mode
    (run-time mask form) at 1e-4 of the cable length against the oracle on the same float32-rounded inputs."""
    n = 16384
    A, B, L, th, ga = _config2_batch(n, 0, sweep_L)
    (o0, o1, o2, o3), Co, vo = orc.transform_catenary_batch(A, B, L, th, ga)
    curves, C, valid = cb.catenary_shapes(A, B, L, th, ga, which=15)
    assert np.array_equal(valid, vo)
    if not sweep_L:
        assert (vo[:n // 100] & 1 == 0).all()               # d > L = 3: taut, no root
    # (slack spans whose C = 2u/l exceeds the reference's bracket end 10 are outside the solve domain too)
    assert 0.4 < (vo & 1).mean() < 0.95 and (vo == 3).mean() > 0.4
    assert_close(C, Co, RTOL64, 0, "C")
    for j, want in enumerate((o0, o1, o2, o3)):
        assert_close(curves[j], want, RTOL64, 1e-12, f"output {j}, all {n} curves")
    # the single-curve compiled forms (W = 1, W = 8) are the same numbers
    (orig,), _, _ = cb.catenary_shapes(A, B, L, th, ga, which=1)
    (fin,), _, _ = cb.catenary_shapes(A, B, L, th, ga, which=8)
    assert np.array_equal(orig, curves[0]) and np.array_equal(fin, curves[3])
    # ---- fp32 fast mode on float32-rounded inputs
    f = np.float32
    A32, B32, L32, th32, ga32 = (x.astype(f) for x in (A, B, L, th, ga))
    (q0, q1, q2, q3), Cq, vq = orc.transform_catenary_batch(A32.astype(np.float64), B32.astype(np.float64),
                                                             L32.astype(np.float64), th32.astype(np.float64),
                                                             ga32.astype(np.float64))
    c32, C32, v32 = cb.catenary_shapes(A32, B32, L32, th32, ga32, which=15, dtype=np.float32)
    agree = v32 == vq
    assert agree.mean() > 0.999, f"validity differs on {(~agree).sum()} of {n}"
    for j, want in enumerate((q0, q1, q2, q3)):
        err = np.abs(c32[j].astype(np.float64) - want)[agree].max(axis=(1, 2)) / L32[agree]
        assert err.max() < RTOL32, f"fp32 output {j}: worst {err.max():.3e} of the cable length"
    both = agree & ((vq & 1) == 1)
    assert np.abs(C32[both].astype(np.float64) / Cq[both] - 1).max() < RTOL32


def test_mpc_float32_controls_under_fp64_arithmetic(cb):
    """The float32 wire format (CAT_MPC_CONTROLS_F32): candidates stored as float, arithmetic fp64.  The oracle on
    the same rounded values is matched at 1e-9 with the exact index -- one launch, the chunked host pipeline (both
    layouts) and the device-resident step."""
    import importlib
    dev = importlib.import_module(cb.__name__ + ".device")
    p = cb.default_mpc_params()
    for K, H in ((3000, 50), (40_037, 12)):
        U32 = _controls(K, H, 21).astype(np.float32)
        pref = cb.circle_reference(H, p["dt"])
        Jo = orc.mpc_rollout_costs(orc.default_mpc_params(), U32.astype(np.float64), pref)
        bi, bc, J = cb.mpc_sample_step(p, U32, pref)                      # a float32 array selects the wire format
        assert J.dtype == np.float64
        assert_close(J, Jo, RTOL64, 0, f"f32-stored controls, K={K}")
        assert (bi, bc) == orc.argmin_mpc(J) and bi == int(np.argmin(Jo))
        Us = np.ascontiguousarray(U32.transpose(1, 2, 0))
        bi2, bc2, J2 = cb.mpc_sample_step(p, Us.astype(np.float64), pref, layout="H3K", controls_dtype=np.float32)
        assert bi2 == bi and bc2 == bc and np.array_equal(J2, J)
        st = dev.MpcDeviceStep(p, pref, K, controls=Us, keep_costs=True)
        st.launch()
        i3, c3 = st.result()
        assert int(i3[0]) == bi and float(c3[0]) == bc and np.array_equal(st.costs.cpu().numpy()[0], J)
        st.close()


def test_mpc_steps_overlapping_on_two_streams(cb):
    """Two device-resident steps that share ONE context, each with a private workspace, launched back to back on two
    CUDA streams (the overlapped throughput loop of bench.py): winners identical to the serial launches."""
    import importlib
    import torch
    dev = importlib.import_module(cb.__name__ + ".device")
    K, H = 20_000, 30
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    ctx = cb.Context(0)
    Us = [np.ascontiguousarray(_controls(K, H, 50 + i).transpose(1, 2, 0)) for i in range(2)]
    steps = [dev.MpcDeviceStep(p, pref, K, controls=Us[i], ctx=ctx, private_workspace=True) for i in range(2)]
    want = []
    for st in steps:
        st.launch()
        i, c = st.result()
        want.append((int(i[0]), float(c[0])))
    streams = [torch.cuda.Stream() for _ in range(2)]
    for rep in range(20):
        for st, s in zip(steps, streams):
            with torch.cuda.stream(s):
                st.launch()
    torch.cuda.synchronize()
    for st, w in zip(steps, want):
        i, c = st.result()
        assert (int(i[0]), float(c[0])) == w
        st.close()
    for k, U in enumerate(Us):
        Jo = orc.mpc_rollout_costs(orc.default_mpc_params(), U.transpose(2, 0, 1), pref)
        assert want[k][0] == int(np.argmin(Jo))


def test_cold_path_counters(cb):
    """cat_mpc_stats: the benign config-3 step never leaves the warm solve; 4x controls at dt = 0.05 does."""
    ctx = cb.default_context()
    K, H = 4096, 50
    p = cb.default_mpc_params()
    ctx.stats(enable=True, reset=True)
    cb.mpc_sample_step(p, _controls(K, H, 3), cb.circle_reference(H, p["dt"]), return_costs=False)
    quiet = ctx.stats(enable=True, reset=True)
    # every warp iterates once, on its first step (the record it starts from is a guess, not a root)
    assert quiet["solve_cold"] <= K // 32 and quiet["sine_cold"] == 0, quiet
    p5 = cb.default_mpc_params(dt=0.05)
    cb.mpc_sample_step(p5, _controls(K, H, 3) * 4.0, cb.circle_reference(H, 0.05), return_costs=False)
    loud = ctx.stats(enable=False, reset=True)
    assert loud["solve_cold"] > 4 * (K // 32) and loud["sine_cold"] > 0, loud
    warp_steps = (K // 32) * H
    assert loud["solve_cold"] <= warp_steps and loud["sine_cold"] <= 4 * warp_steps


def test_back_to_back_steps_overlap_safely(cb):
    """Programmatic dependent launch (cat_set_overlap, default on): a run of steps enqueued on one stream without a
    host sync overlaps -- each winner lands in its own buffer, and the SHARED cost vector, winner slots and tickets
    are touched by a step only after its predecessor has completed.  Same results as with the overlap switched off."""
    import importlib
    import torch
    dev = importlib.import_module(cb.__name__ + ".device")
    K, H = 30_000, 25
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    ctx = cb.Context(0)
    sets = [torch.from_numpy(np.ascontiguousarray(_controls(K, H, 70 + i).transpose(1, 2, 0))).cuda() for i in range(3)]
    st = dev.MpcDeviceStep(p, pref, K, controls=sets[0].cpu().numpy(), keep_costs=True, ctx=ctx)
    ctx.set_overlap(False)
    want, want_costs = [], []
    for U in sets:
        st.set_controls(U)
        st.launch()
        i, c = st.result()
        want.append((int(i[0]), float(c[0])))
        want_costs.append(st.costs.cpu().numpy().copy())
    ctx.set_overlap(True)
    n = 60
    wins = [torch.zeros(2, dtype=torch.int64, device="cuda") for _ in range(n)]
    for j in range(n):
        st.set_controls(sets[j % 3])
        st.launch(winner=wins[j])
    torch.cuda.synchronize()
    for j in range(n):
        w = wins[j].cpu().numpy()
        assert (int(w[1]), float(w[:1].view(np.float64)[0])) == want[j % 3], j
    assert np.array_equal(st.costs.cpu().numpy(), want_costs[(n - 1) % 3])
    # a different kind of work between two steps: plain stream order
    for j in range(6):
        st.set_controls(sets[j % 3])
        st.launch(winner=wins[j])
        wins[j + 10].copy_(wins[j])
    torch.cuda.synchronize()
    for j in range(6):
        assert torch.equal(wins[j + 10], wins[j])
        assert int(wins[j].cpu()[1]) == want[j % 3][0]
    st.close()


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_shapes_every_mask_and_odd_point_counts(cb, dtype):
    """Every output mask (the masks with two or more rotated curves take the two-stores-per-point form of k_shape when
    the point count is even, the others and odd point counts the scalar form): each curve equals, bit for bit, the same
    curve asked for alone."""
    n = 1000 + 37
    A, B, L, th, ga = (x.astype(dtype) for x in _config2_batch(n, 3, True))
    for n_pts in (100, 7, 2):
        alone = {}
        for bit in (1, 2, 4, 8):
            (alone[bit],), C1, v1 = cb.catenary_shapes(A, B, L, th, ga, which=bit, n_pts=n_pts, dtype=dtype)
        for mask in range(1, 16):
            curves, C, valid = cb.catenary_shapes(A, B, L, th, ga, which=mask, n_pts=n_pts, dtype=dtype)
            bits = [b for b in (1, 2, 4, 8) if mask & b]
            assert len(curves) == len(bits)
            for cv, b in zip(curves, bits):
                assert cv.shape == (n, n_pts, 3)
                assert np.array_equal(cv, alone[b], equal_nan=True), (n_pts, mask, b)
