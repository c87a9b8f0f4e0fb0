import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    if os.environ.get("PGPU_PREWARM"):   # tools/prewarm_kernel_cache.py: compile-only operators in every (xdist) worker
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import prewarm_kernel_cache
        prewarm_kernel_cache.patch()


@pytest.fixture(scope="session", autouse=True)
def _native_built():
    # the library must exist; tests never fall back to anything else
    from presto_b200 import build as B
    B.build()
    yield
