"""Shared fixtures. GPU tests are marked `gpu`; everything else must pass on a CPU-only machine."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))  # the oracle is the checker here, never the thing tested as product

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def mod():
    import __graft_entry__ as g
    return g.build()


@pytest.fixture(scope="session")
def syn(mod):
    import importlib
    return importlib.import_module("okvis2x_b200.synthetic")


@pytest.fixture(scope="session")
def ref(mod):
    import refbind
    return refbind.Ref()


@pytest.fixture(scope="session")
def port(mod):
    import refbind
    return refbind.Port()


@pytest.fixture(scope="session")
def ctx(mod):
    c = mod.Context(0)
    yield c
    c.close()


def kp_assert_close(a, b, tol=1e-4):
    """Keypoint parity bar of BASELINE.json: same set and order, response/octave exact, floats within 1e-4."""
    assert len(a) == len(b), (len(a), len(b))
    assert np.array_equal(a["response"], b["response"])
    assert np.array_equal(a["octave"], b["octave"])
    assert np.array_equal(a["class_id"], b["class_id"])
    assert np.array_equal(a["size"], b["size"])
    for f in ("x", "y", "angle"):
        if len(a):
            assert float(np.abs(a[f] - b[f]).max()) <= tol, f
