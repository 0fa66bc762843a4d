"""cat_mpc_step_static_dev_f64: the allocation-free sampling-MPC step (scratch in a fixed device array) against the
existing cat_mpc_step_dev_f64, both called through the C ABI with the same device buffers.  The two run the same
kernel on the same inputs, so costs, winner cost and winner index must agree bit for bit (the fp64 tolerance of the
other tests, 1e-9, is asserted as well).  Collected last on purpose."""
import ctypes as C
import importlib

import numpy as np
import pytest

from conftest import assert_close

pytestmark = pytest.mark.gpu
RTOL64 = 1e-9


def _step(cb, K, H, *, n=1, seed=1234, sampled=False):
    dev = importlib.import_module(cb.__name__ + ".device")
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    sigma = np.array([108.48977412955092, 15.880219723306585, 63.13363738505157])   # scaler.pkl scale_[3:6], mm/s
    if n > 1:
        pref = np.stack([pref + 0.01 * i for i in range(n)])
    if sampled:
        return dev.MpcDeviceStep(p, pref, K, sigma=sigma, seed=seed, keep_costs=True)
    rng = np.random.default_rng(seed)
    U = np.clip(rng.normal(size=(H, 3, K)), -3, 3) * sigma[None, :, None]
    return dev.MpcDeviceStep(p, pref, K, controls=np.ascontiguousarray(U), keep_costs=True)


def _call(st, fn):
    """One call of ``fn`` (an MPC dev-step entry point of the C ABI) on the buffers of ``st``; results on the host."""
    import torch
    st.winner.zero_()
    st.costs.fill_(float("nan"))
    st.ctx.set_stream(torch.cuda.current_stream(st.dev).cuda_stream)
    st.ctx.lib.cat_set_workspace(st.ctx.handle, None)
    st.ctx.check(fn(st.ctx.handle, C.byref(st.params), st.problems.data_ptr(),
                    None if st.controls is None else st.controls.data_ptr(), st.p_ref.data_ptr(),
                    st.winner.data_ptr(), st.winner.data_ptr() + 8 * st.n, st.costs.data_ptr()))
    idx, cost = st.result()
    return idx, cost, st.costs.cpu().numpy()


@pytest.mark.parametrize("sampled", [False, True])
def test_static_step_matches_dev_step_twice_in_a_row(cb, sampled):
    st = _step(cb, 8192 + 37, 50, sampled=sampled)           # ragged: the last block is partly idle
    lib = st.ctx.lib
    i0, c0, J0 = _call(st, lib.cat_mpc_step_dev_f64)
    assert np.isfinite(c0).all() and np.isfinite(J0).all()
    for _ in range(2):                                       # second call: the array holds the first call's contents
        i1, c1, J1 = _call(st, lib.cat_mpc_step_static_dev_f64)
        assert_close(J1, J0, RTOL64, 0, "costs, static scratch vs ctx scratch")
        assert np.array_equal(J1, J0) and np.array_equal(i1, i0) and np.array_equal(c1, c0)


def test_static_step_many_problems_and_other_sizes(cb):
    """Different grid shapes back to back on the shared array (slots and tickets of a wider call must not leak)."""
    for K, H, n in ((4096, 20, 16), (300, 10, 1), (1024, 20, 64)):
        st = _step(cb, K, H, n=n, seed=K, sampled=True)
        i0, c0, J0 = _call(st, st.ctx.lib.cat_mpc_step_dev_f64)
        i1, c1, J1 = _call(st, st.ctx.lib.cat_mpc_step_static_dev_f64)
        assert np.array_equal(J1, J0) and np.array_equal(i1, i0) and np.array_equal(c1, c0)
        assert np.array_equal(i1, np.argmin(J1, axis=1))


def test_static_step_rejects_sizes_beyond_the_array(cb):
    _lib = importlib.import_module(cb.__name__ + "._lib")
    st = _step(cb, 256, 5)
    lib, ctx = st.ctx.lib, st.ctx
    args = (st.problems.data_ptr(), None, st.p_ref.data_ptr(), st.winner.data_ptr(), st.winner.data_ptr() + 8,
            st.costs.data_ptr())
    # more thread blocks than winner slots: nothing is read before the check, so the buffers may stay small
    big = cb.make_params(cb.default_mpc_params(), 128 * _lib.STATIC_MAX_BLOCKS + 1, 5, flags=_lib.MPC_RNG_CONTROLS,
                         sigma=(1.0, 1.0, 1.0))
    assert lib.cat_mpc_step_static_dev_f64(ctx.handle, C.byref(big), *args) == _lib.CAT_EINVAL
    msg = lib.cat_last_error(ctx.handle).decode()
    assert "static scratch" in msg and str(_lib.STATIC_MAX_BLOCKS) in msg
    many = cb.make_params(cb.default_mpc_params(), 128, 5, n_problems=_lib.STATIC_MAX_PROBLEMS + 1,
                          flags=_lib.MPC_RNG_CONTROLS, sigma=(1.0, 1.0, 1.0))
    assert lib.cat_mpc_step_static_dev_f64(ctx.handle, C.byref(many), *args) == _lib.CAT_EINVAL
    # the context is still usable
    i0, c0, J0 = _call(st, lib.cat_mpc_step_dev_f64)
    i1, c1, J1 = _call(st, lib.cat_mpc_step_static_dev_f64)
    assert np.array_equal(J1, J0) and np.array_equal(i1, i0)
