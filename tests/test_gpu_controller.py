"""Receding-horizon controller: the device-resident loop against the oracle's
mpc_receding_horizon on the same seeds.  fp64: 1e-9 relative on the logged trajectory, winner indices exact."""
import numpy as np
import pytest

from conftest import assert_close
from oracle import catenary_oracle as orc

pytestmark = pytest.mark.gpu
SIG = np.array([40.0, 40.0, 20.0])
UMAX = np.array([300.0, 300.0, 300.0])
# columns compared at 1e-9 (13 = step index, 11 = winner index are exact; 15 reserved)
FLOAT_COLS = [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 14]


def _setup(T, H, **over):
    po = orc.default_mpc_params(w_trk=1e3, w_du=1e-6, **over)
    traj = orc.circle_reference(T + H, po["dt"], period=4.0)
    return po, traj


def _check(got, want, rtol=1e-9, atol=1e-12):
    assert np.array_equal(got["log"][:, 11], want["log"][:, 11]), "winner indices differ"
    assert np.array_equal(got["log"][:, 13], want["log"][:, 13])
    assert_close(got["log"][:, FLOAT_COLS], want["log"][:, FLOAT_COLS], rtol, atol, "log")
    assert_close(got["u_nom"], want["u_nom"], rtol, 1e-9, "u_nom")


@pytest.mark.parametrize("use_graph", [True, False])
def test_argmin_loop_matches_oracle(cb, use_graph):
    T, H, K = 25, 20, 768
    po, traj = _setup(T, H)
    want = orc.mpc_receding_horizon(po, traj, T, K, H, SIG, UMAX, seed=5, mode="argmin")
    got = cb.mpc_receding_horizon(cb.default_mpc_params(w_trk=1e3, w_du=1e-6), traj, T, K, H, SIG, UMAX, seed=5,
                                  mode="argmin", use_graph=use_graph)
    assert got["step"] == T
    _check(got, want)
    st = got["state"]
    assert_close(st["P1"], want["state"]["P1"], 1e-9, 0, "final P1")
    assert_close(st["theta"], want["state"]["theta"], 1e-9, 0, "final theta")
    assert_close(st["gamma_prev"], want["state"]["gamma_prev"], 1e-9, 0, "final gamma_prev")
    # the loop tracks: position error stays in the millimetre range while the target moves ~10 cm/s
    err = np.linalg.norm(got["log"][:, :3] - traj[:T], axis=1)
    assert err.max() < 0.01


def test_mppi_loop_matches_oracle(cb):
    T, H, K = 20, 16, 1024
    po, traj = _setup(T, H)
    want = orc.mpc_receding_horizon(po, traj, T, K, H, SIG, UMAX, seed=11, mode="mppi", lam=0.02)
    got = cb.mpc_receding_horizon(cb.default_mpc_params(w_trk=1e3, w_du=1e-6), traj, T, K, H, SIG, UMAX, seed=11,
                                  mode="mppi", lam=0.02)
    _check(got, want)


def test_loop_warm_start_and_reference_hold(cb):
    """Non-zero initial nominal sequence, a reference shorter than the run (last row held), controls that
    hit the u_max clamp."""
    T, H, K = 12, 10, 512
    po, traj = _setup(T, H)
    traj = traj[:8]                                         # t + h >= 8 reads row 7
    u0 = np.tile(np.array([120.0, -80.0, 30.0]), (H, 1))
    umax = np.array([100.0, 100.0, 25.0])
    want = orc.mpc_receding_horizon(po, traj, T, K, H, SIG, umax, seed=3, u_nom0=u0)
    got = cb.mpc_receding_horizon(cb.default_mpc_params(w_trk=1e3, w_du=1e-6), traj, T, K, H, SIG, umax, seed=3,
                                  u_nom0=u0)
    _check(got, want)
    assert np.abs(got["log"][:, 3:6]).max() <= 100.0


def test_loop_generic_equations(cb):
    """A hall-of-fame pair other than the shipped one goes through the generic (switch) kernels."""
    T, H, K = 8, 12, 512
    cx_t, cx_g = 19, 7                                      # hall-of-fame rows (equations_d*_dt.csv)
    po, traj = _setup(T, H, theta_cx=cx_t, gamma_cx=cx_g)
    want = orc.mpc_receding_horizon(po, traj, T, K, H, SIG, UMAX, seed=9)
    got = cb.mpc_receding_horizon(cb.default_mpc_params(w_trk=1e3, w_du=1e-6, theta_cx=cx_t, gamma_cx=cx_g), traj,
                                  T, K, H, SIG, UMAX, seed=9)
    _check(got, want)


def test_controller_object_fleet_and_incremental_runs(cb):
    """Several independent problems in one controller; run() in pieces equals one run; set_state injects a
    measurement; the log window read-back."""
    n, T, H, K = 3, 10, 12, 512
    rng = np.random.default_rng(4)
    poses = np.array([1.0, 1.0, -1.0]) + rng.normal(size=(n, 3)) * 0.1
    p = cb.default_mpc_params(w_trk=1e3, w_du=1e-6)
    prm = dict(p, P1=poses)
    trajs = np.stack([orc.circle_reference(T + H, p["dt"], center=poses[i], period=4.0) for i in range(n)])
    c = cb.RecedingHorizonController(prm, trajs, K, H, SIG, UMAX, seed=21, max_steps=T)
    c.run(4).run(6)
    r = c.read()
    assert r["step"] == T and r["log"].shape == (n, T, 16)
    for i in range(n):
        po = orc.default_mpc_params(w_trk=1e3, w_du=1e-6, P1=tuple(poses[i]))
        want = orc.mpc_receding_horizon(po, trajs[i], T, K, H, SIG, UMAX, seed=21, problem=i)
        _check(dict(log=r["log"][i], u_nom=r["u_nom"][i]), want)
    w = c.read(log_from=3, log_n=4)["log"]
    assert np.array_equal(w, r["log"][:, 3:7])
    with pytest.raises(ValueError):
        c.run(1)                                            # log capacity exhausted
    c.close()
    # measurement injection: after set_state the next step starts from the injected pose
    c = cb.RecedingHorizonController(p, trajs[0], K, H, SIG, UMAX, seed=2, max_steps=4)
    c.run(1)
    s = c.read()["state"].copy()
    s["P1"][0] = [1.05, 0.95, -1.02]
    c.set_state(s)
    c.run(1)
    got = c.read()
    po = orc.default_mpc_params(w_trk=1e3, w_du=1e-6, P1=(1.05, 0.95, -1.02), V1=tuple(s["V1"][0]),
                                A1=tuple(s["A1"][0]), theta=float(s["theta"][0]), gamma=float(s["gamma"][0]),
                                theta_prev=float(s["theta_prev"][0]), gamma_prev=float(s["gamma_prev"][0]))
    U = orc.ctrl_candidates(_unom_after_first(trajs[0], K, H), 2, 1, K, SIG, UMAX)
    win = trajs[0][np.minimum(1 + np.arange(H), len(trajs[0]) - 1)]
    J = orc.mpc_rollout_costs(po, U, win)
    assert int(got["log"][0, 1, 11]) == int(np.argmin(J))
    assert_close(got["log"][0, 1, 10], J.min(), 1e-9, 0, "best cost after set_state")
    c.close()


def _unom_after_first(traj, K, H):
    po = orc.default_mpc_params(w_trk=1e3, w_du=1e-6)
    return orc.mpc_receding_horizon(po, traj, 1, K, H, SIG, UMAX, seed=2)["u_nom"]


def test_loop_fp32_fast_mode(cb):
    """fp32 rollouts + fp64 plant: the closed loop is a chaotic map of its argmin decisions, so the check is
    on behaviour (tracking error, finite log), not on element-wise parity."""
    T, H, K = 25, 20, 2048
    po, traj = _setup(T, H)
    got = cb.mpc_receding_horizon(cb.default_mpc_params(w_trk=1e3, w_du=1e-6), traj, T, K, H, SIG, UMAX, seed=5,
                                  dtype=np.float32)
    assert np.isfinite(got["log"]).all()
    assert np.linalg.norm(got["log"][:, :3] - traj[:T], axis=1).max() < 0.01
    # first step: same candidates up to fp32 rounding -> same winner cost within the fast-mode tolerance
    want = orc.mpc_receding_horizon(po, traj, 1, K, H, SIG, UMAX, seed=5)
    assert_close(got["log"][0, 10], want["log"][0, 10], 1e-4, 0, "first-step best cost (fp32)")


def test_controller_argument_errors(cb):
    p = cb.default_mpc_params()
    traj = orc.circle_reference(30, p["dt"])
    with pytest.raises(ValueError):
        cb.RecedingHorizonController(p, traj, 256, 10, SIG, UMAX, mode="mppi", lam=0.0)
    with pytest.raises(ValueError):
        cb.RecedingHorizonController(p, traj[:, :2], 256, 10, SIG, UMAX)
    with pytest.raises(ValueError):
        cb.RecedingHorizonController(p, traj, 256, 10, SIG, UMAX, mode="cem")
