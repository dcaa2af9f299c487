import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure).  Built on demand with g++."""
    from ldso_b200 import build, capi
    build.build_oracle()
    return capi.Api(build.ORACLE_LIB, "orc_")


@pytest.fixture(scope="session")
def gpu():
    """The product library through its C ABI; no fallback — a missing .so or device is an error."""
    from ldso_b200 import capi
    return capi.load()
