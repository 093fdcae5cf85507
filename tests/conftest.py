import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "reference: needs /root/reference (build container only)")


@pytest.fixture(scope="session", autouse=True)
def _build_oracle():
    """The oracle (test infrastructure) is compiled on demand; the product library is built by
    __graft_entry__.build() and must already exist for the GPU tests."""
    from oracle import slo

    slo.build()
    yield
