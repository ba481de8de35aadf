import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    """The plain-C restatement of the reference CPU path (oracle/rt_oracle.c)."""
    from oracle.loader import PortOracle, build_port
    build_port()
    return PortOracle()


@pytest.fixture(scope="session")
def ref():
    """The compiled reference itself (oracle/_ref); present in the build container and, prebuilt, on the GPU box."""
    from oracle.loader import RefOracle
    if not RefOracle.available():
        pytest.skip("oracle/_ref/libref_oracle.so not built (needs /root/reference)")
    return RefOracle()


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name + ".npz"))
    return load


@pytest.fixture(scope="session")
def be():
    """One rt_ctx on cuda:0 through the C ABI.  No fallback: a missing library or GPU is a hard failure."""
    from raytracinginoneweekend_b200.backend import Backend
    b = Backend(0)
    yield b
    b.close()


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def assert_bits_equal(a, b, what=""):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    if a.dtype == np.float32:
        same = (bits(a) == bits(b)) | (np.isnan(a) & np.isnan(b))
    else:
        same = a == b
    assert same.all(), f"{what}: {int((~same).sum())} of {same.size} elements differ"
