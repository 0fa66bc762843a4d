import os
import sys

import numpy as np
import pytest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built_lib():
    """Build (or reuse) the in-tree shared library; nvcc cross-compiles without a GPU."""
    import __graft_entry__ as ge
    ge.build()
    import catenary_b200 as cb
    assert os.path.exists(cb.LIB_PATH)
    return cb.LIB_PATH


@pytest.fixture(scope="session")
def cb(built_lib):
    import catenary_b200
    return catenary_b200


@pytest.fixture(scope="session")
def ctx(cb):
    return cb.default_context()


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


def assert_close(got, want, rtol, atol=0.0, what=""):
    got = np.asarray(got, np.float64)
    want = np.asarray(want, np.float64)
    assert got.shape == want.shape, f"{what}: shape {got.shape} != {want.shape}"
    assert np.array_equal(np.isnan(got), np.isnan(want)), f"{what}: NaN pattern differs"
    m = ~np.isnan(want)
    err = np.abs(got[m] - want[m])
    tol = atol + rtol * np.abs(want[m])
    bad = err > tol
    assert not bad.any(), f"{what}: {bad.sum()} of {m.sum()} outside tol; worst abs {err.max():.3e}"
