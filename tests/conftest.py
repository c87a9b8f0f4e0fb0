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


# GPU cases added since the last visit to a B200 (profiles/r01g_pytest_gpu.log) run after everything that has already
# passed there, so that under `-x` a surprise in a new case cannot hide the state of the proven ones.
NEWEST_GPU_CASES = ("nullif_grid_", "simple_case_", "case_nullif_into_aggregation", "optest_", "test_gpu_parquet", "test_gpu_operators", "test_gpu_intermediate")


def pytest_collection_modifyitems(config, items):
    def newest(item):
        return item.get_closest_marker("gpu") is not None and any(k in item.nodeid for k in NEWEST_GPU_CASES)
    items.sort(key=newest)  # stable: everything else keeps its order
