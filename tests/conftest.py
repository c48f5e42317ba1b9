import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ctx():
    """One stitch_ctx on cuda:0 for the GPU tests (fails loudly without a GPU)."""
    from image_stitch_b200 import _native
    c = _native.Context(int(os.environ.get("STITCH_B200_DEVICE", "0")))
    yield c
    c.close()
