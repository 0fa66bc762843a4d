"""CPU-side checks: the C-ABI library builds, loads and exports every symbol include/*.h declares; the
code generator reproduces the oracle's lambdified equations; host-only helpers behave."""
import ctypes
import importlib
import math
import os
import re

import numpy as np
import pytest

from conftest import ROOT
from oracle import catenary_oracle as orc

PKG = "catenary-model-estimation-and-mpc-control-for-rov-tethered-systems_b200"


def _declared():
    with open(os.path.join(ROOT, "include", "catenary_b200.h")) as fh:
        return sorted(set(re.findall(r"CAT_API[^;(]*?\b(cat_\w+)\s*\(", fh.read())))


def test_library_exports_every_declared_symbol(cb):
    names = _declared()
    assert len(names) >= 38
    lib = ctypes.CDLL(cb.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/catenary_b200.h but not exported"
    bound = importlib.import_module(PKG + "._lib").SIGNATURES
    assert sorted(bound) == names, "ctypes SIGNATURES and the header disagree"
    assert lib.cat_abi_version() == 3


def test_struct_layouts_match_header(cb):
    L = importlib.import_module(PKG + "._lib")
    assert ctypes.sizeof(L.CatMpcProblem) == 16 * 8
    assert ctypes.sizeof(L.CatMpcParams) == 16 + 6 * 4 + 13 * 8 + 8 + 3 * 8 + 8
    assert L.CatMpcParams.dt.offset == 40 and L.CatMpcParams.seed.offset == 144


def test_problem_records_are_the_c_struct(cb):
    """make_problems fills CatMpcProblem through a flat (n,16) view: byte-identical to per-field assignment, scalar
    and per-problem values broadcast alike."""
    lib = importlib.import_module(cb.__name__ + "._lib")
    rng = np.random.default_rng(3)
    p = cb.default_mpc_params()
    for n, prm in ((1, p), (5, dict(p, P1=rng.normal(size=(5, 3)), V1=rng.normal(size=(5, 3)), theta=rng.normal(size=5),
                                    gamma_prev=rng.normal(size=5)))):
        got = cb.make_problems(prm, n)
        want = np.zeros(n, lib.PROBLEM_DTYPE)
        for k in ("P0", "P1", "V1", "A1"):
            want[k] = np.broadcast_to(np.asarray(prm[k], np.float64), (n, 3))
        for k in ("theta", "gamma", "theta_prev", "gamma_prev"):
            want[k] = np.broadcast_to(np.asarray(prm[k], np.float64), (n,))
        assert got.dtype == lib.PROBLEM_DTYPE and got.shape == (n,) and got.flags["C_CONTIGUOUS"]
        assert got.tobytes() == want.tobytes()


def test_no_cpu_fallback_without_device(cb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(cb.CatenaryLibraryError):
        cb.Context(0)
    with pytest.raises(cb.CatenaryLibraryError):
        cb.solve_catenary(1.0, 0.0, 3.0)


def test_reduce_winners_order(cb):
    inf, nan = float("inf"), float("nan")
    assert cb.reduce_winners([3.0, 1.0, 1.0, 2.0], [10, 30, 20, 5]) == (20, 1.0)       # tie -> lowest index
    assert cb.reduce_winners([3.0, nan, 1.0], [0, 7, 9]) == (9, 1.0)                   # a NaN cost never wins ...
    i, c = cb.reduce_winners([nan, inf, nan], [8, 6, 3])
    assert i == 3 and math.isnan(c)                                                     # ... it ranks with +inf
    i, c = cb.reduce_winners([nan, nan], [5, 4])
    assert i == 4 and math.isnan(c)                                                     # all NaN: lowest index
    assert cb.reduce_winners([inf, inf], [4, 2]) == (2, inf)
    rng = np.random.default_rng(0)
    c = rng.integers(0, 5, 64).astype(float)
    assert cb.reduce_winners(c, np.arange(64))[0] == int(np.argmin(c))


def test_codegen_matches_sympy_lambdify():
    cg = importlib.import_module(PKG + ".codegen")
    rng = np.random.default_rng(3)
    X = rng.normal(size=(50, 18))
    table = cg.equation_table()
    assert len(table) == 37
    for eq_id, name, cx, s in table:
        node = cg.parse(s)
        want = orc.equation(name, cx)(X)
        got = np.array([cg.evaluate(node, list(x), math) for x in X])
        assert np.allclose(got, want, rtol=1e-13, atol=1e-15), (name, cx)
    ids = cg.shipped_ids()
    assert table[ids["dtheta_dt"]][2] == 13 and table[ids["dgamma_dt"]][2] == 3      # eq_*.txt
    with open(cg.OUT_FILE) as fh:
        txt = fh.read()
    assert "cat_eq_dtheta_dt_c13" in txt and "CAT_N_EQUATIONS 37" in txt


def test_codegen_rejects_unknown_syntax():
    cg = importlib.import_module(PKG + ".codegen")
    for bad in ("x1 + foo(x2)", "x1 ** x2", "x1 +", "2 $ 3"):
        with pytest.raises(cg.ParseError):
            cg.parse(bad)
    assert cg.live_features(cg.parse("sin(x3)*x17 - Abs(x2)**2")) == {2, 3, 17}


def test_scaler_constants_in_artifacts():
    m = orc.load_model()
    assert m["n_features"] == 18 and len(m["scaler_mean"]) == 18
    assert abs(m["scaler_mean"][3] - 80.85390753943355) < 1e-12
    assert abs(m["scaler_scale"][17] - 0.017328840568644522) < 1e-15


def test_product_never_imports_oracle():
    pkg_dir = os.path.join(ROOT, PKG)
    for dp, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dp, f)) as fh:
                    src = fh.read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f


def test_sampler_table_definition_and_header():
    """The tabulated inverse normal CDF that DEFINES the device sampler: the oracle's evaluation against
    scipy.special.ndtri, the first moments of the sampled controls, and the generated CUDA header carrying exactly
    the numbers of artifacts/icdf_table.json."""
    from scipy.special import ndtri
    from oracle import catenary_oracle as orc
    e_lo, m, coef = orc.icdf_table()
    assert (e_lo, m) == (5, 5) and coef.shape == (26 * 32, 4)
    rng = np.random.default_rng(0)
    r = rng.integers(0, 2 ** 32, size=400_000, dtype=np.uint64).astype(np.uint32)
    r[:8] = [0, 1, 31, 32, 2 ** 31 - 1, 2 ** 31, 2 ** 32 - 33, 2 ** 32 - 1]            # both tails and the centre
    z = orc.icdf_normal(r)
    rp = np.maximum(np.where(r >= 2 ** 31, ~r, r).astype(np.int64), 32)               # the definition clamps r' at 32
    want = np.where(r >= 2 ** 31, 1.0, -1.0) * -ndtri((rp + 0.5) / 2.0 ** 32)
    assert np.abs(z - want).max() < 1e-9
    assert z[5] > 0 > z[4] and abs(z[4] + z[5]) < 1e-12                                # antisymmetric about 2^31
    assert abs(z.mean()) < 0.01 and abs(z.std() - 1) < 0.01
    U = orc.sample_controls(77, np.arange(20_000), 12, [108.0, 16.0, 63.0])
    assert np.abs(U).max(axis=(0, 1)).tolist() == [324.0, 48.0, 189.0]                # clipped at 3 sigma exactly
    assert np.allclose(U.std(axis=(0, 1)) / [108.0, 16.0, 63.0], 0.9975, atol=0.01)   # std of a normal clipped at 3 sigma
    cg = importlib.import_module(PKG + ".codegen")
    text = open(cg.generate_icdf()).read()
    nums = [float(x) for x in text.split("{", 1)[1].split("}", 1)[0].replace("\n", " ").split(",") if x.strip()]
    assert np.array_equal(np.array(nums).reshape(-1, 4), coef)
