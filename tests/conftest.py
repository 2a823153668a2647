import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def frb():
    """The product package with the CUDA library loaded.  Fails loudly (never skips) if libfrb.so is missing."""
    import fractal_renderer_b200 as f

    f._lib.load()
    return f


@pytest.fixture(scope="session")
def gpu(frb):
    """A visible B200.  GPU tests must not pass on a silent fallback: no device => failure, not skip."""
    n = frb.device_count()
    assert n >= 1, "no CUDA device visible: -m gpu tests need a B200 (there is no CPU path to fall back to)"
    frb.init(0)
    return frb
