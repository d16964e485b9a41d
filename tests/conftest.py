import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the checkers exist (CPU tier builds them; the GPU box gets the
    prebuilt .so files with the snapshot but can rebuild the C oracle)."""
    import oracle
    if not oracle.available("port"):
        oracle.build(ref=False, port=True)
    if not oracle.available("ref") and os.path.isdir(oracle.REFERENCE_SRC):
        oracle.build(ref=True, port=False)
    yield
