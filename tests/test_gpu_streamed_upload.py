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


def _ctx(cb, streamed):
    import importlib
    Context = importlib.import_module(cb.__name__ + ".context").Context
    old = os.environ.get("CAT_MPC_STREAMED")
    os.environ["CAT_MPC_STREAMED"] = "1" if streamed else "0"     # read once, in cat_init
    try:
        return Context(0)
    finally:
        if old is None:
            del os.environ["CAT_MPC_STREAMED"]
        else:
            os.environ["CAT_MPC_STREAMED"] = old


def _controls(K, H, seed):
    rng = np.random.default_rng(seed)
    return np.ascontiguousarray((np.clip(rng.normal(size=(H, 3, K)), -3, 3) * SIGMA[None, :, None]))


@pytest.mark.parametrize("K,H", [(65536, 50), (40064, 50), (40000 + 17, 50), (32768, 9), (32768, 8), (36000, 7), (33024, 1),
                                 (75776, 3)])
def test_streamed_equals_chunked(cb, K, H):
    U = _controls(K, H, seed=K + H)
    p = cb.default_mpc_params()
    pref = cb.circle_reference(H, p["dt"])
    c_str, c_chk = _ctx(cb, True), _ctx(cb, False)
    for kw in (dict(), dict(controls_dtype=np.float32), dict(dtype=np.float32)):
        Uk = U.astype(np.float32) if kw.get("dtype") is np.float32 else U
        want = cb.mpc_sample_step(p, Uk, pref, layout="H3K", ctx=c_chk, **kw)
        for _ in range(2):                                    # second call: next epoch on the same arrival words
            got = cb.mpc_sample_step(p, Uk, pref, layout="H3K", ctx=c_str, **kw)
            assert got[0] == want[0] and got[1] == want[1], (K, H, kw)
            assert np.array_equal(got[2], want[2]), (K, H, kw)


def test_streamed_config3_vs_oracle(cb):
    K, H = 65536, 50
    U = _controls(K, H, seed=1234)
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
