"""Host MPC step with the streamed upload (ONE rollout launched at once, candidates following over PCIe in row groups
with arrival words) against the chunked upload (CAT_MPC_STREAMED=0: eight chunk launches + a fold) and the oracle.
Same kernel arithmetic on the same values, so costs and winner agree bit for bit between the two uploads."""
import os

import numpy as np
import pytest

from conftest import assert_close
from oracle import catenary_oracle as orc

pytestmark = pytest.mark.gpu
SIGMA = np.array(orc.load_model()["scaler_scale"][3:6])


def _ctx(cb, streamed, **env):
    """A context whose library switches (read once, in cat_init) are set for this context only."""
    import importlib
    Context = importlib.import_module(cb.__name__ + ".context").Context
    env = dict({"CAT_MPC_STREAMED": "1" if streamed else "0"}, **env)
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        return Context(0)
    finally:
        for k, v in old.items():
            if v is None:
                del os.environ[k]
            else:
                os.environ[k] = v


_PIN = []


def _controls(K, H, seed, cb=None):
    """(H, 3, K) candidates; page-locked when ``cb`` is given (the streamed upload takes page-locked memory only --
    small ordinary arrays, which the Context does not lock on sight, go up chunked)."""
    rng = np.random.default_rng(seed)
    U = np.clip(rng.normal(size=(H, 3, K)), -3, 3) * SIGMA[None, :, None]
    if cb is None:
        return np.ascontiguousarray(U)
    if not _PIN:
        _PIN.append(_ctx(cb, True))            # owner of the page-locked buffers of this module
    out = _PIN[0].pinned_empty((H, 3, K))
    out[:] = U
    return out


@pytest.mark.parametrize("K,H", [(65536, 50), (40064, 50), (40000 + 17, 50), (32768, 9), (32768, 8), (36000, 7), (33024, 1),
                                 (75776, 3)])
def test_streamed_equals_chunked(cb, K, H):
    U = _controls(K, H, seed=K + H, cb=cb)
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    c_str, c_chk = _ctx(cb, True), _ctx(cb, False)
    U32 = _PIN[0].pinned_empty(U.shape, np.float32)
    U32[:] = U
    for kw in (dict(), dict(controls_dtype=np.float32), dict(dtype=np.float32)):
        Uk = U32 if kw else U
        want = cb.mpc_sample_step(p, Uk, pref, layout="H3K", ctx=c_chk, **kw)
        for _ in range(2):                                    # second call: next epoch on the same arrival words
            got = cb.mpc_sample_step(p, Uk, pref, layout="H3K", ctx=c_str, **kw)
            assert got[0] == want[0] and got[1] == want[1], (K, H, kw)
            assert np.array_equal(got[2], want[2]), (K, H, kw)


def test_streamed_config3_vs_oracle(cb):
    K, H = 65536, 50
    U = _controls(K, H, seed=1234, cb=cb)
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    bi, bc, J = cb.mpc_sample_step(p, U, pref, layout="H3K", ctx=_ctx(cb, True))
    pick = np.random.default_rng(0).choice(K, 4096, replace=False)
    Jo = orc.mpc_rollout_costs(orc.default_mpc_params(), np.ascontiguousarray(U[:, :, pick].transpose(2, 0, 1)), pref)
    assert_close(J[pick], Jo, 1e-9, 0, "streamed upload, costs vs oracle")
    assert bi == int(np.argmin(J)) and bc == J[bi]
    # without the costs array, and alternating with a device-sampled call on the same context
    ctx = _ctx(cb, True)
    for _ in range(3):
        b2, c2, none = cb.mpc_sample_step(p, U, pref, layout="H3K", ctx=ctx, return_costs=False)
        assert none is None and b2 == bi and c2 == bc
        cb.mpc_sample_step(p, None, pref, sigma=SIGMA, K=4096, ctx=ctx)


def test_streamed_many_groups_and_copied_arrival_words(cb):
    """Eight row groups (the most the kernel takes) and arrival words written by 4-byte copies instead of stream
    memory operations: the fallback for drivers without them."""
    K, H = 65536, 41
    U = _controls(K, H, seed=7, cb=cb)
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    want = cb.mpc_sample_step(p, U, pref, layout="H3K", ctx=_ctx(cb, False))
    for env in (dict(CAT_MPC_STREAM_SPLIT="10,10,10,5,3,1,1,1"), dict(CAT_MPC_STREAM_MEMOP="0"),
                dict(CAT_MPC_STREAM_SPLIT="40,1", CAT_MPC_STREAM_MEMOP="0")):
        ctx = _ctx(cb, True, **env)
        for _ in range(2):
            got = cb.mpc_sample_step(p, U, pref, layout="H3K", ctx=ctx)
            assert got[0] == want[0] and got[1] == want[1] and np.array_equal(got[2], want[2]), env


def test_streamed_repeated_calls_alternating_shapes(cb):
    """Sixty calls on one context, two shapes and two wire types in turn: arrival words, epochs and the control
    buffer are reused call after call."""
    p = cb.default_mpc_params()
    cases = []
    for K, H, kw in ((65536, 50, dict(controls_dtype=np.float32)), (36864, 12, dict())):
        U = _controls(K, H, seed=K, cb=cb)
        if kw:
            U32 = _PIN[0].pinned_empty(U.shape, np.float32)
            U32[:] = U
            U = U32
        pref = cb.circle_reference(H, p["dt"])
        want = cb.mpc_sample_step(p, U, pref, layout="H3K", ctx=_ctx(cb, False), return_costs=False, **kw)
        cases.append((U, pref, kw, want))
    ctx = _ctx(cb, True)
    for i in range(60):
        U, pref, kw, want = cases[i % 2]
        got = cb.mpc_sample_step(p, U, pref, layout="H3K", ctx=ctx, return_costs=False, **kw)
        assert got[0] == want[0] and got[1] == want[1], i


def test_ordinary_arrays_take_the_chunked_upload_and_agree(cb):
    """Candidates in ordinary (pageable) memory -- too small for the Context to page-lock on sight -- and odd sizes
    go up chunked; same results as from page-locked memory."""
    p = cb.default_mpc_params()
    for K, H in ((33024, 1), (40000 + 17, 6)):
        U = _controls(K, H, seed=K)
        pref = cb.circle_reference(H, p["dt"])
        a = cb.mpc_sample_step(p, U, pref, layout="H3K", ctx=_ctx(cb, True))
        b = cb.mpc_sample_step(p, _controls(K, H, seed=K, cb=cb), pref, layout="H3K", ctx=_ctx(cb, True))
        assert a[0] == b[0] and a[1] == b[1] and np.array_equal(a[2], b[2])
