import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "animating-gaussian-splats_b200")
for p in (ROOT, PKG, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    return oracle_lib.load()


@pytest.fixture(scope="session")
def refdgr():
    """The UNMODIFIED reference extension built by oracle/Makefile into oracle/_ref (Oracle A)."""
    import oracle_lib
    mod = oracle_lib.load_reference_ext()
    if mod is None:
        pytest.skip("oracle/_ref/refdgr/_C.so not built (make -C oracle ref)")
    return mod
