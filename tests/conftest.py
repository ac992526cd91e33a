import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_sessionstart(session):
    """Builds libncdrift.so / libncdhost.so (nvcc cross-compiles without a GPU) and the CPU oracle
    when they are missing or older than their sources, so the suite runs from a fresh checkout."""
    from nexus_configuration_controller_b200 import build as b

    b.build_native(force=False)
    import oracle

    if not os.path.exists(oracle.LIB_PATH) or any(
        os.path.getmtime(os.path.join(ROOT, "oracle", f)) > os.path.getmtime(oracle.LIB_PATH)
        for f in os.listdir(os.path.join(ROOT, "oracle")) if f.endswith((".cpp", ".h"))
    ):
        oracle.build()


@pytest.fixture(scope="session")
def ctx():
    """One GPU context for the whole session (gpu tests only)."""
    from nexus_configuration_controller_b200.drift import Context

    c = Context(0)
    yield c
    c.close()
