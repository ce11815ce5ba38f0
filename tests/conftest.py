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
    """The reference oracle (oracle/_ref, the unmodified reference code compiled here)."""
    from oracle import ref as r
    if not r.available():
        pytest.skip("oracle/_ref/libusgscsm_ref.so not built")
    return r
