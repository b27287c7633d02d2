"""Shared fixtures. `-m gpu` tests need a B200 and call the product through its C-ABI;
everything else runs on CPU. The oracle (oracle/) is the checker only."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference compiled into oracle/_ref/libref_oracle.so (travels to the GPU box prebuilt)."""
    from oracle import refo
    try:
        return refo.RefOracle()
    except (FileNotFoundError, OSError) as e:
        pytest.skip(f"reference oracle not built: {e}")


@pytest.fixture(scope="session")
def b200():
    import libsnark_b200 as m
    m.init(0)
    return m
