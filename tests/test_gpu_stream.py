"""Streaming runs (DESIGN.md "streaming runs"; csrc/device/pgpu_device.cuh body_runs_count / body_runs_write / body_runs_emit /
body_runs_to_table): a long device-resident page whose group keys arrive in key order is aggregated run by run into a linear
buffer, without a hash table; a key order that breaks later sends the buffered partial states into the table. Which path took
the rows is never observable: every case equals the CPU oracle bit for bit (quantities are integer-valued DOUBLEs, sums are
exact in any order). The thresholds that keep short pages and small tables away from the path are lowered with PGPU_TUNE_STREAM_*
knobs so the cases stay small."""
import numpy as np
import pytest

from presto_b200 import tpch
from presto_b200.operators import Aggregation, GpuPipelineOperator, plan_text, toPages
from presto_b200.relational import compare, constant, field
from presto_b200.spi import BIGINT, DOUBLE, INTEGER, IntArrayBlock, LongArrayBlock, Page
from test_gpu_sweep import NO_DIRECTORY, DevicePageOf
from util import assert_rows_equal_ignore_order, rows_of, run_oracle


def run_streamed(plan, pages, monkeypatch, expect_streamed=None, host_pages=(), stream=1, resets=0, force=0):
    monkeypatch.setenv("PGPU_TUNE_STREAM", str(stream))
    monkeypatch.setenv("PGPU_TUNE_STREAM_MIN_ROWS", "1")
    monkeypatch.setenv("PGPU_TUNE_STREAM_MIN_TABLE_MB", "0")
    monkeypatch.setenv("PGPU_TUNE_STREAM_FORCE", str(force))   # 1: whatever the probe of the first keys says
    monkeypatch.setenv("PGPU_TUNE_SWEEP", "0")
    dps = [DevicePageOf(pg) for pg in pages]
    op = GpuPipelineOperator(plan, 0)
    try:
        for _ in range(resets + 1):
            for dp in dps:
                op.addDeviceInput(dp.cols, dp.ncols, dp.rows)
            mid = op.stats()
            out = rows_of(toPages(op, list(host_pages)))
            st = op.stats()
            if _ < resets:
                op.reset()
    finally:
        op.close()
        for dp in dps:
            dp.free()
    if expect_streamed is not None:
        assert (mid.streamed_runs > 0) == expect_streamed, mid.streamed_runs
    return out, st


def sorted_page(n, seed=1, max_run=7, key_nulls=False, start=-20_000):
    """keys ascending (as unsigned words: from `start` >= 0 on; negative starts cross the sign once) in runs of 1..max_run rows"""
    rng = np.random.default_rng(seed)
    lens = rng.integers(1, max_run + 1, n)
    k = np.repeat(start + np.cumsum(rng.integers(1, 9, n)), lens)[:n].astype(np.int64)
    q = rng.integers(1, 51, n).astype(np.float64)
    qnull = (rng.random(n) < 0.1).astype(np.uint8)
    knull = None
    if key_nulls:  # NULL keys in one block at the end: one more run
        knull = np.zeros(n, np.uint8)
        knull[-17:] = 1
    return Page(LongArrayBlock(BIGINT, k, knull), LongArrayBlock(DOUBLE, q, qnull))


def sum_plan(filter=None, aggs=None):
    aggs = aggs or [Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT), Aggregation("max", DOUBLE, 1, DOUBLE)]
    projections = None if filter is None else [field(0, BIGINT), field(1, DOUBLE)]
    return plan_text([BIGINT, DOUBLE], filter, projections, step="single", groupByChannels=[0], aggregations=aggs, expectedGroups=NO_DIRECTORY)


@pytest.mark.gpu
@pytest.mark.parametrize("rows", [1, 5, 127, 128, 129, 1023, 1024, 1025, 4 * 1024 * 37 + 11, 400_003])
def test_stream_sorted_keys(rows, monkeypatch):
    plan, page = sum_plan(), sorted_page(rows, start=1)
    act, st = run_streamed(plan, [page], monkeypatch, expect_streamed=True)
    assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=1, rtol=0.0)


@pytest.mark.gpu
def test_stream_lineitem_orderkey(monkeypatch):
    """the bench's GROUP BY l_orderkey shape: the generator's keys ascend in runs of 1..7"""
    plan, page = tpch.highcard_plan(NO_DIRECTORY), tpch.highcard_page(0, 300_007, 300_007, False)
    act, st = run_streamed(plan, [page], monkeypatch, expect_streamed=True)
    exp = run_oracle(plan, [page])
    assert st.streamed_runs >= len(exp) and st.streamed_runs < len(exp) + 300_007 // 128 + 2   # one extra row per run a warp boundary cut
    assert_rows_equal_ignore_order(act, exp, nkeys=1, rtol=0.0)


@pytest.mark.gpu
def test_stream_off_by_knob_and_for_unordered_keys(monkeypatch):
    plan, page = sum_plan(), sorted_page(100_000, start=1)
    act, st = run_streamed(plan, [page], monkeypatch, expect_streamed=False, stream=0)
    assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=1, rtol=0.0)
    plan, page = tpch.highcard_plan(NO_DIRECTORY), tpch.highcard_page(0, 100_000, 100_000, True)
    act, st = run_streamed(plan, [page], monkeypatch, expect_streamed=False)
    assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=1, rtol=0.0)


@pytest.mark.gpu
@pytest.mark.parametrize("where", ["late", "sign", "second_page", "repeat_page"])
def test_stream_order_breaks_after_the_probe(where, monkeypatch):
    """the probe sees the first 65 536 keys in order; the order breaks later: the emit kernel counts the decrease and the
    buffered partial states go through the hash table. Keys that occur on both sides of the break must merge."""
    if where == "late":
        page = sorted_page(200_000, start=1)
        page.blocks[0].values[150_000:] = page.blocks[0].values[:50_000]          # the first 50 000 rows' keys come back
        pages = [page]
    elif where == "sign":
        pages = [sorted_page(200_000, start=-100_000)]                           # ascending as signed, one descent as unsigned words
        assert pages[0].blocks[0].values[0] < 0 < pages[0].blocks[0].values[-1]
    elif where == "second_page":
        pages = [sorted_page(100_000, seed=3, start=1_000_000), sorted_page(100_000, seed=4, start=1)]
    else:
        pages = [sorted_page(100_000, seed=3, start=1)] * 2                       # every key twice, 100 000 rows apart
    plan = sum_plan()
    act, st = run_streamed(plan, pages, monkeypatch, expect_streamed=True)
    assert_rows_equal_ignore_order(act, run_oracle(plan, pages), nkeys=1, rtol=0.0)


@pytest.mark.gpu
def test_stream_pages_that_continue_in_order(monkeypatch):
    """three device pages, the key order carries on across them (the last key of a page = the first of the next: one group)"""
    whole = sorted_page(300_000, start=1)
    cuts = [0, 99_999, 200_001, 300_000]
    pages = [Page(*[type(b)(b.type, b.values[a:z].view(np.float64 if b.type == DOUBLE else np.int64),
                            None if b.valueIsNull is None else b.valueIsNull[a:z]) for b in whole.blocks]) for a, z in zip(cuts, cuts[1:])]
    plan = sum_plan()
    act, st = run_streamed(plan, pages, monkeypatch, expect_streamed=True)
    assert_rows_equal_ignore_order(act, run_oracle(plan, [whole]), nkeys=1, rtol=0.0)


@pytest.mark.gpu
def test_stream_then_host_pages(monkeypatch):
    """host pages after the device page: the buffered runs move into the table, the host pages aggregate on top"""
    dev = sorted_page(150_000, seed=5, start=1)
    host = [sorted_page(3_000, seed=6, start=1), sorted_page(2_000, seed=7, start=500_000)]
    plan = sum_plan()
    act, st = run_streamed(plan, [dev], monkeypatch, expect_streamed=True, host_pages=host)
    assert_rows_equal_ignore_order(act, run_oracle(plan, [dev] + host), nkeys=1, rtol=0.0)


@pytest.mark.gpu
def test_stream_filter_null_keys_and_long_groups(monkeypatch):
    """rows the filter rejects leave dead runs between live ones (a key's rows on both sides still merge); NULL keys form a
    group; a group far longer than a warp's 128 rows is cut into pieces that the emit kernel puts together again; one longer than
    64 pieces goes through the table"""
    flt = compare("lt", field(1, DOUBLE), constant(40.0, DOUBLE))
    for max_run, start in ((7, 1), (700, 1), (30_000, 1)):
        page = sorted_page(250_000, seed=max_run, max_run=max_run, key_nulls=True, start=start)
        plan = sum_plan(filter=flt)
        act, st = run_streamed(plan, [page], monkeypatch, expect_streamed=True, force=1)
        assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=1, rtol=0.0)


@pytest.mark.gpu
def test_stream_signed_keys_nulls_last_and_sorted_two_field_keys(monkeypatch):
    """the expected order is the one a sorted table comes in (P::key_order): signed integers through zero, NULLs after the values,
    the first grouping field first. None of these falls back to the table (no rehash, no table at all)."""
    plan = sum_plan()
    page = sorted_page(200_000, start=-100_000, key_nulls=True)
    act, st = run_streamed(plan, [page], monkeypatch, expect_streamed=True)
    assert st.table_capacity == 0, "fell back to the table"
    assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=1, rtol=0.0)
    # (orderkey, linenumber): sorted by both, every row its own group, the second field restarts at 1 in every order
    rng = np.random.default_rng(5)
    n = 150_000
    lens = rng.integers(1, 8, n)
    k1 = np.repeat(np.cumsum(rng.integers(1, 5, n)) - 70_000, lens)[:n].astype(np.int64)
    k2 = (np.arange(n) - np.repeat(np.concatenate([[0], np.cumsum(lens)[:-1]]), lens)[:n] + 1).astype(np.int32)
    q = rng.integers(1, 51, n).astype(np.float64)
    page = Page(LongArrayBlock(BIGINT, k1), IntArrayBlock(INTEGER, k2), LongArrayBlock(DOUBLE, q))
    plan = plan_text([BIGINT, INTEGER, DOUBLE], step="single", groupByChannels=[0, 1],
                     aggregations=[Aggregation("sum", DOUBLE, 2, DOUBLE), Aggregation("count", BIGINT)], expectedGroups=NO_DIRECTORY)
    act, st = run_streamed(plan, [page], monkeypatch, expect_streamed=True)
    assert st.table_capacity == 0, "fell back to the table"
    exp = run_oracle(plan, [page])
    assert len(exp) == n
    assert_rows_equal_ignore_order(act, exp, nkeys=2, rtol=0.0)


@pytest.mark.gpu
def test_stream_two_key_fields_and_int_keys(monkeypatch):
    rng = np.random.default_rng(11)
    n = 180_000
    k1 = np.repeat(np.arange(1, n + 1), rng.integers(1, 40, n))[:n].astype(np.int32)       # ascending, long runs
    k2 = (np.arange(n) // 3 % 5).astype(np.int64)                                            # changes inside a k1 run, not in order
    q = rng.integers(1, 51, n).astype(np.float64)
    page = Page(IntArrayBlock(INTEGER, k1), LongArrayBlock(BIGINT, k2), LongArrayBlock(DOUBLE, q))
    plan = plan_text([INTEGER, BIGINT, DOUBLE], step="single", groupByChannels=[0, 1],
                     aggregations=[Aggregation("sum", DOUBLE, 2, DOUBLE), Aggregation("min", DOUBLE, 2, DOUBLE)], expectedGroups=NO_DIRECTORY)
    act, st = run_streamed(plan, [page], monkeypatch)
    assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=2, rtol=0.0)


@pytest.mark.gpu
def test_stream_reset_and_rerun(monkeypatch):
    plan, page = sum_plan(), sorted_page(120_000, seed=9, start=1)
    act, st = run_streamed(plan, [page], monkeypatch, expect_streamed=True, resets=2)
    assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=1, rtol=0.0)
