import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    oracle_lib.lib()
    return oracle_lib


@pytest.fixture(scope="session")
def nb():
    """The ctypes binding of libnbody_gpu.so (built on demand; never falls back to the CPU)."""
    import pnb200
    mod = pnb200.load()
    if not os.path.exists(mod.LIB_PATH):
        mod.build()
    mod.lib()
    return mod


@pytest.fixture(scope="session")
def ppm(tmp_path_factory):
    import oracle_lib
    p = str(tmp_path_factory.mktemp("ppm") / "nbody.ppm")
    oracle_lib.write_ppm(p)
    return p
