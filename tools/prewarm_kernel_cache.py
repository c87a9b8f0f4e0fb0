#!/usr/bin/env python
"""Fill the in-tree kernel cache (presto_b200/_kernel_cache, travels to the GPU box with the snapshot) without a GPU.

The plan kernels are generated per plan and page signature and NVRTC-compiled for sm_100a on first use (~2 s each); the
cache key covers the device library, so every edit of csrc/device/pgpu_device.cuh invalidates all of them and the next
GPU test run would spend a minute of box time compiling. This script runs the `-m gpu` tests with GpuPipelineOperator
patched to compile-only pipelines (device -1): addInput compiles the variant of the page's signature instead of running
it, outputs are empty. Test OUTCOMES ARE MEANINGLESS here (most tests fail their assertions) and are discarded; only the
cache matters. Also compiles the bench's plans under the tuning knobs named on the command line:

    python tools/prewarm_kernel_cache.py [PGPU_TUNE_RUNS=1,PGPU_TUNE_LOCB=4 ...]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from presto_b200 import _native as N  # noqa: E402
from presto_b200 import operators as O  # noqa: E402
from presto_b200 import tpch  # noqa: E402

_init = O.GpuPipelineOperator.__init__


def _signature(cols, n):
    enc, nullable = [], []
    for i in range(n):
        c = cols[i]
        enc.append(c.encoding)
        if c.encoding in (N.ENC_DICTIONARY, N.ENC_RLE):
            nullable.append(bool(c.dictionary) and bool(c.dictionary.contents.nulls))
        else:
            nullable.append(bool(c.nulls))
    return enc, nullable


def patch():
    def init(self, plan, device=0):
        _init(self, plan, -1)
        self._done = False

    def add_input(self, page, pinned=False):
        cols, keep = self._marshal(page.blocks)
        self.compileSignature(*_signature(cols, len(page.blocks)))

    def add_device_input(self, cols, ncols, rows):
        self.compileSignature(*_signature(cols, ncols))

    def finish(self):
        self._done = True

    cls = O.GpuPipelineOperator
    cls.__init__ = init
    cls.addInput = add_input
    cls.addDeviceInput = add_device_input
    cls.addSerializedInput = lambda self, payload, n: None
    cls.addParquetRowGroup = lambda self, chunks, rows: None
    cls.getOutput = lambda self: None
    cls.getSerializedOutput = lambda self: None
    cls.needsInput = lambda self: not self._done
    cls.finish = finish
    cls.isFinished = lambda self: self._done
    cls.sync = lambda self: None
    cls.reset = lambda self: None


def bench_plans():
    rows = 600_037_902
    ops = [O.GpuPipelineOperator(p, -1) for p in (tpch.q1_plan(), tpch.q1_plan("partial"), tpch.q6_plan(), tpch.q6_plan("partial"),
                                                  tpch.highcard_plan(rows // 4), tpch.highcard_plan(rows // 4, "partial"),
                                                  tpch.highcard_plan(16_000_000 // 4), tpch.highcard_plan(64_000_000 // 4))]
    ops[0].compileSignature([1, 1, 1, 0, 1, 1, 0], [False] * 7)
    ops[1].compileSignature([1, 1, 1, 0, 1, 1, 0], [False] * 7)
    for op in ops:
        op.close()
    gather_variants()


def gather_variants():
    """FINAL plans as the gather exchange feeds them (pgpu_exchange_partial): every channel with a null vector, VARCHAR keys
    as 16-byte cells; plus the device-page signature of the repartition shape"""
    import bench as B
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_oracle_protocol import q1_final_plan
    from test_gpu_exchange import FINAL_HC, REPARTITION
    plans = [B.final_plan_for({"name": "q1"}, None), B.final_plan_for({"name": "q6"}, None), B.final_plan_for({"name": "highcard", "rows": 600_037_902}, None),
             q1_final_plan(), FINAL_HC(REPARTITION), FINAL_HC(1000)]
    for plan in plans:
        from presto_b200.exchange import gather_signature
        op = O.GpuPipelineOperator(plan, -1)
        op.compileSignature(*gather_signature(plan))
        op.close()


if __name__ == "__main__":
    # start from an empty cache: whatever is in it afterwards is exactly what the current sources compile to (stale cubins of
    # earlier device-library versions would only travel to the GPU box as dead weight)
    import shutil
    shutil.rmtree(os.path.join(ROOT, "presto_b200", "_kernel_cache"), ignore_errors=True)
    before = len(os.listdir(os.path.join(ROOT, "presto_b200", "_kernel_cache"))) if os.path.isdir(os.path.join(ROOT, "presto_b200", "_kernel_cache")) else 0
    for spec in ["", "PGPU_TUNE_BUCKETS=1"] + sys.argv[1:]:
        env = dict(kv.split("=") for kv in spec.split(",") if kv)
        os.environ.update(env)
        bench_plans()
        for k in env:
            os.environ.pop(k)
    import pytest
    patch()
    # the patch has to reach xdist workers too: they import this module's patch() through conftest's PGPU_PREWARM hook
    os.environ["PGPU_PREWARM"] = "1"
    pytest.main([os.path.join(ROOT, "tests"), "-m", "gpu", "-q", "--timeout", "300", "-p", "no:cacheprovider", "--no-header", "--tb=no",
                 "-n", str(min(8, os.cpu_count() or 1))])
    after = len(os.listdir(os.path.join(ROOT, "presto_b200", "_kernel_cache")))
    print("kernel cache: %d -> %d files" % (before, after))
