"""Rows A8/A9 of SURVEY.md 8a on the GPU: compute_catenary_3D (models/catenary_3d.py:5-39) and the planar
level-support profile (catenary_model.py:10-31) against reference-generated vectors and the oracle."""
import numpy as np
import pytest

from conftest import assert_close, golden
from oracle import catenary_oracle as orc

pytestmark = pytest.mark.gpu


def _atol(p0, p1, L):
    """The reference's fixed point diverges (oracle.catenary_3d_parameter): its own sag values carry an absolute
    cancellation error of ~a * 2^-52, so that is the floor of any comparison."""
    with np.errstate(all="ignore"):
        a = abs(orc.catenary_3d_parameter(p0, p1, float(L)))
    return 1e-12 + (32 * 2.0 ** -52 * a if np.isfinite(a) and a < 1e12 else 0.0)


def test_compute_catenary_3d_golden_and_batch(cb):
    g = golden("catenary_3d.npz")
    for i in range(len(g["L"])):
        got = cb.compute_catenary_3D(g["p0"][i], g["p1"][i], float(g["L"][i]), 100)
        assert got.shape == (100, 3)
        assert_close(got, g["ref_pts"][i], 1e-9, _atol(g["p0"][i], g["p1"][i], g["L"][i]), f"catenary_3d golden {i}")
    rng = np.random.default_rng(5)
    n = 4000
    p0 = rng.normal(size=(n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    p1 = p0 + d * rng.uniform(0.2, 3.3, n)[:, None]          # ~9 % taut (distance >= 3)
    L = np.where(rng.random(n) < 0.5, 3.0, rng.uniform(2.0, 5.0, n))
    got = cb.compute_catenary_3D(p0, p1, L, 64)
    assert got.shape == (n, 64, 3)
    for i in list(range(0, n, 97)) + list(np.flatnonzero(np.linalg.norm(p1 - p0, axis=1) >= L)[:5]):
        with np.errstate(all="ignore"):
            want = orc.compute_catenary_3D(p0[i], p1[i], float(L[i]), 64)
        assert_close(got[i], want, 1e-9, _atol(p0[i], p1[i], L[i]), f"rope {i}")
    one = cb.compute_catenary_3D(p0, p1, 3.0, 10)            # scalar length broadcast
    assert_close(one[3], orc.compute_catenary_3D(p0[3], p1[3], 3.0, 10), 1e-9, _atol(p0[3], p1[3], 3.0), "scalar L")
    assert np.array_equal(cb.compute_catenary_3D(p0[0], p0[0] + [4.0, 0, 0], 3.0, 5),
                          np.linspace(p0[0], p0[0] + [4.0, 0, 0], 5))


def test_planar_profile_functions(cb):
    g = golden("catenary_model.npz")
    for j, C in enumerate(g["eq_C"]):
        # the reference forms cosh(Cx) - 1 with cancellation: absolute error ~1e-16 / C near x = 0
        assert_close(cb.catenary_equation(g["eq_x"], float(C)), g["ref_eq"][j], 1e-9, 2e-15 / C, f"equation C={C}")
    assert_close(cb.catenary_equation(g["eq_x"], np.full(41, 0.7)), g["ref_eq"][1], 1e-9, 1e-14, "per-sample C")
    for (L, l), C_ref, x_ref, y_ref in zip(g["cases"], g["ref_C"], g["ref_x"], g["ref_y"]):
        C = cb.solve_catenary_params(float(L), float(l), 0.0)
        if C_ref > 9.99:                                     # no root below 10: the reference's minimiser hugs the bound
            assert C is None
            continue
        # the reference minimises the squared residual with xatol 1e-5; this is the root itself
        assert abs(C - C_ref) < 2e-5 * max(1.0, C_ref)
        assert abs(2 * np.sinh(C * l / 2) / C - L) < 1e-12 * L
        x, y = cb.generate_catenary(None, float(L), float(l), 0.0)
        assert np.array_equal(x, x_ref)
        assert_close(y, y_ref, 1e-4, 1e-6, "profile (reference C is only 1e-5 accurate)")
        assert_close(y, orc_profile(x, C), 1e-9, 1e-14, "profile at the same C")
    with pytest.raises(ValueError):
        cb.generate_catenary(None, 5.0, 0.5, 0.0)


def orc_profile(x, C):
    return (np.cosh(C * x) - 1) / C
