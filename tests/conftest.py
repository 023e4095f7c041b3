import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (runs on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Builds libcrt.so / liboracle.so (and oracle/_ref where the reference is mounted) if they are missing."""
    from cudaraytracer_b200 import capi
    import oracle_lib

    if not (os.path.exists(capi.LIB_PATH) and os.path.exists(oracle_lib.ORACLE_SO) and
            os.path.exists(os.path.join(ROOT, "examples", "headless_render"))):
        import __graft_entry__

        __graft_entry__.build()
    yield
