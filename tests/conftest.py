import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    oracle_lib.build_core()
    return oracle_lib


@pytest.fixture(scope="session")
def capi():
    """The C-ABI library; built on demand (nvcc cross-compiles without a GPU)."""
    from lz77_sss_b200 import build, capi as c
    if not os.path.exists(c.LIB_PATH):
        build.build()
    return c
