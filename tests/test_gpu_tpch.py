"""GPU parity on the BASELINE.json configs at sizes the oracle finishes in seconds: CUDA path (through the C ABI)
vs the CPU oracle on identical synthetic lineitem columns. Bit-exact for keys/counts, 1e-12 relative for DOUBLE."""
import ctypes

import numpy as np
import pytest

from presto_b200 import _native as N
from presto_b200 import tpch
from presto_b200.operators import GpuPipelineOperator, toPages
from presto_b200.spi import Page
from util import (assert_rows_equal_ignore_order, exact_q1, exact_q6, rows_of, run_gpu, run_oracle, sequential_bound)

pytestmark = pytest.mark.gpu


def split(page_fn, total, sizes):
    pages, r0 = [], 0
    for n in sizes:
        pages.append(page_fn(r0, n, total))
        r0 += n
    assert r0 == total
    return pages


def test_q6_host_pages():
    total = 600_011
    pages = split(tpch.q6_page, total, [200_000, 1, 300_003, 100_007])
    plan = tpch.q6_plan()
    exp = run_oracle(plan, pages)
    act = run_gpu(plan, pages)
    assert len(exp) == 1 and exp[0][0] is not None
    exact, nsel = exact_q6(pages)
    assert_rows_equal_ignore_order(act, exact)                                   # GPU: 1e-12 of the correctly rounded sum
    assert_rows_equal_ignore_order(exp, exact, rtol=sequential_bound(nsel))      # oracle: sequential-order bound


def test_q1_host_pages():
    total = 500_003
    pages = split(tpch.q1_page, total, [123_457, 376_546])
    plan = tpch.q1_plan()
    exp = run_oracle(plan, pages)
    act = run_gpu(plan, pages)
    assert len(exp) == 4
    exact = exact_q1(pages)
    assert_rows_equal_ignore_order(act, exact, nkeys=2)                                   # GPU vs correctly rounded sums
    assert_rows_equal_ignore_order(exp, exact, nkeys=2, rtol=sequential_bound(total))    # oracle (sequential adds)
    # counts are bit-exact
    assert sorted(r[-1] for r in act) == sorted(r[-1] for r in exp)


def test_q1_dictionary_pages_match_plain():
    total = 300_001
    plain = split(tpch.q1_page, total, [100_000, 200_001])
    enc = split(tpch.q1_page_encoded, total, [100_000, 200_001])
    plan = tpch.q1_plan()
    exp = run_oracle(plan, plain)
    assert run_oracle(plan, enc) == exp                       # same values, same order: bit-identical on the CPU
    assert_rows_equal_ignore_order(run_gpu(plan, enc), exact_q1(plain), nkeys=2)
    assert_rows_equal_ignore_order(run_gpu(plan, plain), exact_q1(plain), nkeys=2)


@pytest.mark.parametrize("shuffled", [False, True])
def test_highcard_host_pages(shuffled):
    total = 400_000
    pages = split(lambda a, b, c: tpch.highcard_page(a, b, c, shuffled), total, [150_000, 250_000])
    plan = tpch.highcard_plan(expectedGroups=1000)  # deliberately too small: the table must grow
    exp = run_oracle(plan, pages)
    act = run_gpu(plan, pages)
    assert len(exp) > 50_000
    assert_rows_equal_ignore_order(act, exp, nkeys=1)


def _device_result(plan, spec, rows):
    dp = tpch.DevicePage(0, spec, rows)
    op = GpuPipelineOperator(plan, 0)
    try:
        op.addDeviceInput(dp.cols, len(spec), rows)
        op.finish()
        out = []
        while not op.isFinished():
            p = op.getOutput()
            if p is not None:
                out.append(p)
        return rows_of(out), op.stats()
    finally:
        op.close()
        dp.free()


def test_q1_device_page_equals_host_page():
    rows = 1_000_003
    plan = tpch.q1_plan()
    act, st = _device_result(plan, tpch.Q1_DEVICE_SPEC, rows)
    pages = [tpch.q1_page(0, rows, rows)]
    assert_rows_equal_ignore_order(act, exact_q1(pages), nkeys=2)
    exp = run_oracle(plan, pages)
    assert sorted(r[-1] for r in act) == sorted(r[-1] for r in exp)
    assert st.kernel_launches >= 1 and st.h2d_bytes == 0


def test_q6_device_page():
    rows = 2_000_001
    plan = tpch.q6_plan()
    act, _ = _device_result(plan, tpch.Q6_DEVICE_SPEC, rows)
    exact, nsel = exact_q6([tpch.q6_page(0, rows, rows)])
    assert nsel > 10_000
    assert_rows_equal_ignore_order(act, exact)


def test_q6_sum_is_deterministic_run_to_run():
    rows = 1_000_000
    plan = tpch.q6_plan()
    a, _ = _device_result(plan, tpch.Q6_DEVICE_SPEC, rows)
    b, _ = _device_result(plan, tpch.Q6_DEVICE_SPEC, rows)
    assert a == b  # fixed-order block reduction


def test_highcard_device_page_large():
    # size-independent property at a larger size: per-group sums add up to the global sum and the group count is
    # the number of distinct order keys of the generator (7 orders per 28 rows)
    rows = 28 * 400_000
    plan = tpch.highcard_plan(expectedGroups=rows // 4)
    act, st = _device_result(plan, tpch.HC_DEVICE_SPEC, rows)
    assert len(act) == rows // 4
    total = sum(r[1] for r in act)
    qty = tpch.host_column("quantity", 0, rows, rows)
    assert abs(total - float(qty.sum())) <= 1e-9 * float(qty.sum())
    assert st.table_rehashes == 0
