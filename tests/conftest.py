import os
import sys

import pytest

# the CPU checkers use OpenMP; on over-committed vCPUs a spinning barrier is pathologically slow
os.environ.setdefault("OMP_WAIT_POLICY", "passive")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: long-running known-answer runs")


@pytest.fixture(scope="session")
def lib():
    """The built C-ABI library (compiled on demand; nvcc cross-compiles without a GPU)."""
    from libgran_b200 import build, system

    build.build()
    return system.load_library()


@pytest.fixture(scope="session")
def oracle():
    from oracle import bindings

    bindings.build_oracle()
    return bindings
