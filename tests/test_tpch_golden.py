"""Parity on REAL TPC-H rows, pinned by the reference's own fixtures (tests/golden/README.md):

* the row generator — the oracle's sequential restatement (oracle/tpch_dbgen.cpp) and the product's random-access one
  (presto_b200/csrc/tpch_dbgen.cuh, host and device) — against the table statistics the reference ships for its tpch
  connector at sf0.01 and sf1 (row counts, min / max / distinct values of every column on the path) and against each other;
* TPC-H Q1 and Q6 at sf1 through the oracle (CPU tests) and through the CUDA path (-m gpu) against the reference's golden
  answers hive_tpch/q01.result and q06.result: bit-exact on the DECIMAL(12,2) schema those answers were computed on, and
  within 1e-9 relative on the tpch connector's DOUBLE schema (the golden file is decimal arithmetic; DOUBLE-vs-DOUBLE parity
  is held to 1e-12 in test_gpu_tpch.py)."""
import ctypes
import json
import os
from decimal import Decimal

import numpy as np
import pytest

import oracle_op
from presto_b200 import _native as N
from presto_b200 import tpch
from util import assert_rows_equal_ignore_order, exact_q1, run_gpu, run_oracle, sequential_bound

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
STATS = json.load(open(os.path.join(GOLDEN, "tpch_lineitem_stats.json")))
ORACLE_COLUMNS = [("orderkey", np.int64), ("linenumber", np.int32), ("quantity", np.int64), ("extendedprice", np.int64), ("discount", np.int64),
                  ("tax", np.int64), ("shipdate", np.int32), ("commitdate", np.int32), ("receiptdate", np.int32), ("returnflag", np.uint8),
                  ("linestatus", np.uint8), ("partkey", np.int64)]
Q1_SCALES = (None, None, 2, 2, 4, 6, 2, 2, 2, 0)  # output scales of hive_tpch/q01.sql on DECIMAL(12,2) columns
PAGE_ORDERS = 250_000


def oracle_lineitem(sf, first_order=0, n_orders=None):
    """rows of orders [first_order, first_order + n_orders) from the oracle's generator, as integer columns"""
    h = oracle_op.lib()
    h.oracle_tpch_order_count.restype = ctypes.c_int64
    h.oracle_tpch_order_count.argtypes = [ctypes.c_double]
    h.oracle_tpch_lineitem.restype = ctypes.c_int64
    h.oracle_tpch_lineitem.argtypes = [ctypes.c_double, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64] + [ctypes.c_void_p] * len(ORACLE_COLUMNS)
    if n_orders is None:
        n_orders = h.oracle_tpch_order_count(sf) - first_order
    cols = {k: np.empty(n_orders * 7, dt) for k, dt in ORACLE_COLUMNS}
    n = h.oracle_tpch_lineitem(sf, first_order, n_orders, n_orders * 7, *[c.ctypes.data for c in cols.values()])
    assert n >= 0
    return {k: v[:n] for k, v in cols.items()}


def golden_rows(name, scales):
    rows = []
    for line in open(os.path.join(GOLDEN, name)):
        if line.startswith("--") or not line.strip():
            continue
        cells = line.rstrip("\n").split("|")[:len(scales)]
        rows.append(tuple(c if s is None else int(Decimal(c).scaleb(s)) for c, s in zip(cells, scales)))
    return rows


# ------------------------------------------------------------------------------------------------- the generator
@pytest.mark.parametrize("sf_name,sf", [("sf0.01", 0.01), ("sf1.0", 1.0)])
def test_generator_matches_reference_table_statistics(sf_name, sf):
    want = STATS[sf_name]
    c = oracle_lineitem(sf)
    assert len(c["orderkey"]) == want["rowCount"] == tpch.tpch_rows(sf)
    as_stat = {"l_extendedprice": lambda v: v / 100.0, "l_discount": lambda v: v / 100.0, "l_tax": lambda v: v / 100.0,
               "l_returnflag": chr, "l_linestatus": chr}
    for name, w in want["columns"].items():
        v = c[name[2:]]
        f = as_stat.get(name, lambda x: x)
        assert (f(v.min()), f(v.max()), len(np.unique(v))) == (w["min"], w["max"], w["distinctValuesCount"]), name


@pytest.mark.parametrize("sf,first,n", [(0.01, 0, 15000), (1.0, 0, 300_000), (1.0, 1_234_567, 65_537), (100.0, 149_999_000, 1000), (1000.0, 1_499_990_000, 4097)])
def test_product_generator_equals_oracle_generator(sf, first, n):
    """random access by order (two modular powers per order) == stepping every stream from the first order"""
    c = oracle_lineitem(sf, first, n)
    rows = len(c["orderkey"])
    assert tpch.tpch_rows(sf, first, n) == rows
    for name in ("orderkey", "shipdate", "returnflag", "linestatus"):
        assert np.array_equal(tpch.tpch_host_column(name, sf, first, n), c[name]), name
    for name in ("quantity", "extendedprice", "discount", "tax"):
        scale = 100 if name == "quantity" else 1
        assert np.array_equal(tpch.tpch_host_column(name, sf, first, n, decimal=True), c[name] * scale), name
        assert np.array_equal(tpch.tpch_host_column(name, sf, first, n), c[name] * scale / 100.0), name   # TpchRecordSet: hundredths / 100.0
    assert np.array_equal(tpch.tpch_host_column("quantity_id", sf, first, n), c["quantity"] - 1)
    assert np.array_equal(tpch.tpch_host_column("discount_id", sf, first, n), c["discount"])
    assert np.array_equal(tpch.tpch_host_column("tax_id", sf, first, n), c["tax"])
    assert np.array_equal(np.array([65, 78, 82], dtype=np.uint8)[tpch.tpch_host_column("returnflag_id", sf, first, n)], c["returnflag"])
    assert np.array_equal(np.array([70, 79], dtype=np.uint8)[tpch.tpch_host_column("linestatus_id", sf, first, n)], c["linestatus"])


def test_generator_argument_errors():
    L = N.lib()
    r = ctypes.c_int64()
    assert L.pgpu_tpch_lineitem_rows(1.0, 0, 1_500_001, ctypes.byref(r)) == N.INVALID_ARGUMENT     # past the last order
    assert L.pgpu_tpch_lineitem_rows(1.0, -1, 10, ctypes.byref(r)) == N.INVALID_ARGUMENT
    buf = np.empty(8, np.int64)
    assert L.pgpu_tpch_lineitem_host(1.0, 0, 100, N.LI["orderkey"], 0, 8, buf.ctypes.data, ctypes.byref(r)) == N.INVALID_ARGUMENT  # too small
    assert r.value == tpch.tpch_rows(1.0, 0, 100)
    assert L.pgpu_tpch_lineitem_host(1.0, 0, 1, N.LI["shipdate"], 1, 8, buf.ctypes.data, None) == N.INVALID_ARGUMENT   # no decimal dates
    assert L.pgpu_tpch_lineitem_host(1.0, 0, 1, N.LI["orderkey_shuffled"], 0, 8, buf.ctypes.data, None) == N.INVALID_ARGUMENT


# -------------------------------------------------------------------------------------- Q1 / Q6 at sf1 vs golden
def sf1_pages(make, decimal):
    total = tpch.tpch_order_count(1.0)
    return [make(1.0, o, min(PAGE_ORDERS, total - o), decimal) for o in range(0, total, PAGE_ORDERS)]


def check_q1_decimal(rows):
    assert sorted(rows) == golden_rows("hive_tpch_q01.result", Q1_SCALES)


def check_q1_double(rows):
    want = [tuple(v if s is None else v / 10.0 ** s for v, s in zip(r, Q1_SCALES)) for r in golden_rows("hive_tpch_q01.result", Q1_SCALES)]
    rows = sorted(rows)
    assert [r[:2] + r[9:] for r in rows] == [(r[0], r[1], int(r[9])) for r in want]          # keys and counts exactly
    for got, exp in zip(rows, want):
        for c in range(2, 6):
            assert abs(got[c] - exp[c]) <= 1e-9 * abs(exp[c]), (got, exp, c)
        for c in range(6, 9):                                                                # golden averages are rounded to cents
            assert abs(got[c] - exp[c]) <= 0.005 + 1e-9, (got, exp, c)


def test_oracle_q1_sf1_decimal_equals_golden_answer():
    check_q1_decimal(run_oracle(tpch.q1_decimal_plan(), sf1_pages(tpch.tpch_q1_page, True)))


def test_oracle_q6_sf1_decimal_equals_golden_answer():
    assert run_oracle(tpch.q6_decimal_plan(), sf1_pages(tpch.tpch_q6_page, True)) == golden_rows("hive_tpch_q06.result", (4,))


def test_oracle_q1_q6_sf1_double_schema_agrees_with_golden_answer():
    check_q1_double(run_oracle(tpch.q1_plan(), sf1_pages(tpch.tpch_q1_page, False)))
    (got,), = run_oracle(tpch.q6_plan(), sf1_pages(tpch.tpch_q6_page, False))
    assert abs(got - 123141078.2283) <= 1e-9 * 123141078.2283


@pytest.mark.gpu
def test_gpu_q1_sf1_decimal_equals_golden_answer():
    check_q1_decimal(run_gpu(tpch.q1_decimal_plan(), sf1_pages(tpch.tpch_q1_page, True)))


@pytest.mark.gpu
def test_gpu_q6_sf1_decimal_equals_golden_answer():
    assert run_gpu(tpch.q6_decimal_plan(), sf1_pages(tpch.tpch_q6_page, True)) == golden_rows("hive_tpch_q06.result", (4,))


@pytest.mark.gpu
def test_gpu_q1_q6_sf1_double_schema():
    pages = sf1_pages(tpch.tpch_q1_page, False)
    rows = run_gpu(tpch.q1_plan(), pages)
    check_q1_double(rows)
    assert_rows_equal_ignore_order(rows, exact_q1(pages), nkeys=2, rtol=1e-12)                       # correctly rounded sums
    assert_rows_equal_ignore_order(rows, run_oracle(tpch.q1_plan(), pages), nkeys=2, rtol=sequential_bound(6_001_215))  # the oracle adds in row order
    (got,), = run_gpu(tpch.q6_plan(), sf1_pages(tpch.tpch_q6_page, False))
    assert abs(got - 123141078.2283) <= 1e-9 * 123141078.2283


# ------------------------------------------------------------------------------------------- device generator
@pytest.mark.gpu
@pytest.mark.parametrize("sf,first,n", [(0.01, 0, 15000), (1.0, 1_234_567, 65_537), (100.0, 149_000_000, 300_001), (1000.0, 1_499_990_000, 4097)])
def test_device_generator_equals_host_generator(sf, first, n):
    L = N.lib()
    rows = tpch.tpch_rows(sf, first, n)
    dev = L.pgpu_device_alloc(0, rows * 8 + 64)
    try:
        cases = [(name, False) for name in ("orderkey", "quantity", "extendedprice", "discount", "tax", "shipdate", "returnflag", "linestatus",
                                            "discount_id", "tax_id", "quantity_id", "returnflag_id", "linestatus_id")]
        cases += [(name, True) for name in ("quantity", "extendedprice", "discount", "tax")]
        for name, decimal in cases:
            want = tpch.tpch_host_column(name, sf, first, n, decimal, rows)
            got = np.empty_like(want)
            r = ctypes.c_int64()
            assert L.pgpu_tpch_lineitem_device(0, sf, first, n, N.LI[name], 1 if decimal else 0, rows, dev, ctypes.byref(r)) == 0
            assert r.value == rows
            assert L.pgpu_memcpy_d2h(0, got.ctypes.data, dev, got.nbytes) == 0
            assert np.array_equal(got, want), (name, decimal)
        r = ctypes.c_int64()
        assert L.pgpu_tpch_lineitem_device(0, sf, first, n, N.LI["orderkey"], 0, rows - 1, dev, ctypes.byref(r)) == N.INVALID_ARGUMENT
        assert r.value == rows
    finally:
        L.pgpu_device_free(0, dev)


def run_device_page(plan, spec, dbgen):
    from presto_b200.operators import GpuPipelineOperator
    from util import rows_of
    dp = tpch.DevicePage(0, spec, dbgen=dbgen)
    op = GpuPipelineOperator(plan, 0)
    try:
        op.addDeviceInput(dp.cols, len(spec), dp.rows)
        op.finish()
        out = []
        while not op.isFinished():
            p = op.getOutput()
            if p is not None:
                out.append(p)
        return rows_of(out)
    finally:
        op.close()
        dp.free()


@pytest.mark.gpu
def test_gpu_device_resident_sf1_equals_golden_answers():
    """the whole sf1 table generated in HBM and aggregated in one pass (the bench's --data dbgen path)"""
    sf1 = (1.0, 0, tpch.tpch_order_count(1.0))
    check_q1_decimal(run_device_page(tpch.q1_decimal_plan(), tpch.Q1_DECIMAL_DEVICE_SPEC, sf1))
    assert run_device_page(tpch.q6_decimal_plan(), tpch.Q6_DECIMAL_DEVICE_SPEC, sf1) == golden_rows("hive_tpch_q06.result", (4,))
    check_q1_double(run_device_page(tpch.q1_plan(), tpch.Q1_DEVICE_SPEC, sf1))       # the staged Q1 kernel on real rows
    (got,), = run_device_page(tpch.q6_plan(), tpch.Q6_DEVICE_SPEC, sf1)
    assert abs(got - 123141078.2283) <= 1e-9 * 123141078.2283
    groups = run_device_page(tpch.highcard_plan(1_500_000), tpch.HC_DEVICE_SPEC, sf1)   # GROUP BY l_orderkey, sum(l_quantity)
    assert len(groups) == 1_500_000
    assert sum(g[1] for g in groups) == 37734107 + 991417 + 74476040 + 37719753 + sum_quantity_after_cutoff()


def sum_quantity_after_cutoff():
    c = oracle_lineitem(1.0)
    return int(c["quantity"][c["shipdate"] > 10471].sum())
