import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o

    o.lib()
    return o


@pytest.fixture(scope="session")
def cuda_lib():
    """The built C-ABI library; GPU tests fail loudly when it is missing (no fallback)."""
    import torch

    import parallax_b200

    assert torch.cuda.is_available(), "GPU test without a CUDA device"
    if not os.path.exists(parallax_b200._lib.LIB_PATH):
        parallax_b200.build()
    return parallax_b200._lib.lib()


@pytest.fixture()
def exact_trig(oracle):
    """Switch the oracle to the engine's sin/cos/atan2 definitions for exact-equality tests."""
    oracle.set_trig_mode(1)
    yield
    oracle.set_trig_mode(0)


@pytest.fixture(scope="session")
def cuda_lib_cpu():
    """The built C-ABI library for host-only entry points (no device needed, no compute calls)."""
    import parallax_b200

    if not os.path.exists(parallax_b200._lib.LIB_PATH):
        parallax_b200.build()
    return parallax_b200._lib.lib()
