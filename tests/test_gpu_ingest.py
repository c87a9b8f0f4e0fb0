"""The ingest path at the reference's page shape: streams of small host pages (RecordPageSource hands out <= 4096-row pages,
spi/connector/RecordPageSource.java:84-131; PageProcessor <= 8192-row batches, operator/project/PageProcessor.java:56-57) are
coalesced into super-pages inside pgpu_add_page / pgpu_add_serialized_page. Whatever way the rows are cut into pages, block
encodings and staging slots, the result must be the one the oracle computes page by page."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import pages_serde as S  # noqa: E402

from presto_b200 import tpch  # noqa: E402
from presto_b200.operators import Aggregation, GpuPipelineOperator, plan_text, toPages  # noqa: E402
from presto_b200.spi import (BIGINT, DOUBLE, DictionaryBlock, LongArrayBlock, Page, RunLengthEncodedBlock, VariableWidthBlock,  # noqa: E402
                             createVarcharType)
from util import assert_rows_equal_ignore_order, exact_q1, exact_q6, rows_of, run_oracle, sequential_bound  # noqa: E402

pytestmark = pytest.mark.gpu


def cut(page_fn, total, sizes):
    pages, r0 = [], 0
    for n in sizes:
        pages.append(page_fn(r0, n, total))
        r0 += n
    assert r0 == total
    return pages


def random_sizes(total, lo, hi, seed):
    rng = np.random.default_rng(seed)
    out = []
    while total > 0:
        n = int(min(total, rng.integers(lo, hi + 1)))
        out.append(n)
        total -= n
    return out


def run_gpu_stats(plan, pages, serialized=False):
    op = GpuPipelineOperator(plan, 0)
    try:
        if serialized:
            for p in pages:
                assert op.needsInput()
                op.addSerializedInput(S.write_raw_page(p.blocks), p.position_count)
            op.finish()
            out = []
            while not op.isFinished():
                q = op.getOutput()
                if q is not None and q.position_count:
                    out.append(q)
        else:
            out = toPages(op, pages)
        return rows_of(out), op.stats()
    finally:
        op.close()


def test_small_pages_are_one_launch():
    total = 300_007
    sizes = random_sizes(total, 1, 4096, 1)
    pages = cut(tpch.q1_page, total, sizes)
    plan = tpch.q1_plan()
    act, st = run_gpu_stats(plan, pages)
    assert_rows_equal_ignore_order(act, exact_q1(pages), nkeys=2)
    exp = run_oracle(plan, pages)
    assert sorted(r[-1] for r in act) == sorted(r[-1] for r in exp)
    assert st.input_pages == len(sizes) and st.input_rows == total
    assert st.super_pages == 1                      # ~150 pages, one super-page
    assert st.kernel_launches <= 4                  # staged tiles + ragged tail + emit
    assert st.h2d_bytes == total * 46


def test_q6_small_pages():
    total = 250_001
    pages = cut(tpch.q6_page, total, random_sizes(total, 1, 8192, 2))
    act, st = run_gpu_stats(tpch.q6_plan(), pages)
    exact, nsel = exact_q6(pages)
    assert_rows_equal_ignore_order(act, exact)
    assert st.super_pages == 1


def test_pages_larger_than_a_staging_slot(monkeypatch):
    monkeypatch.setenv("PGPU_STAGE_MB", "1")        # the smallest slots: one staged tile per SM (284 160 Q1 rows)
    total = 700_003
    pages = cut(tpch.q1_page, total, [350_000, 1, 99_999, 250_003])
    plan = tpch.q1_plan()
    act, st = run_gpu_stats(plan, pages)
    assert_rows_equal_ignore_order(act, exact_q1(pages), nkeys=2)
    assert st.super_pages >= 3
    hc = cut(tpch.highcard_page, total, [350_000, 1, 99_999, 250_003])
    hplan = tpch.highcard_plan(expectedGroups=1000)  # grows while super-pages are in flight
    assert_rows_equal_ignore_order(run_gpu_stats(hplan, hc)[0], run_oracle(hplan, hc), nkeys=1)


def test_signature_changes_flush_the_super_page():
    total = 90_000
    a = cut(tpch.q1_page, total, [10_000] * 9)
    b = cut(tpch.q1_page_encoded, total, [10_000] * 9)
    pages = [a[0], a[1], b[2], b[3], a[4], b[5], b[6], b[7], a[8]]   # plain, plain, dict, dict, plain, dict x 3, plain
    # a null vector on one page (no NULLs in it, but the block carries the array) is another signature
    blk = pages[4].blocks
    pages[4] = Page(blk[0], blk[1], LongArrayBlock(DOUBLE, blk[2].values.view(np.float64), np.zeros(10_000, dtype=np.uint8)), *blk[3:])
    plan = tpch.q1_plan()
    act, st = run_gpu_stats(plan, pages)
    assert_rows_equal_ignore_order(act, exact_q1(a), nkeys=2)
    assert st.super_pages == 5                      # [plain x2] [dict x2] [plain+nulls] [dict x3] [plain]


def test_dictionary_change_flushes_the_super_page():
    # the same ids under another dictionary must not be appended to a super-page that carries the first one
    n = 5000
    ids = np.arange(n, dtype=np.int32) % 3
    keys = LongArrayBlock(BIGINT, np.arange(n, dtype=np.int64) % 7)
    d1 = LongArrayBlock(DOUBLE, np.array([1.0, 2.0, 3.0]))
    d2 = LongArrayBlock(DOUBLE, np.array([10.0, 20.0, 30.0]))
    d1again = LongArrayBlock(DOUBLE, np.array([1.0, 2.0, 3.0]))         # equal content, another object: may be coalesced
    pages = [Page(keys, DictionaryBlock(d1, ids)), Page(keys, DictionaryBlock(d1again, ids)), Page(keys, DictionaryBlock(d2, ids)),
             Page(keys, RunLengthEncodedBlock(LongArrayBlock(DOUBLE, np.array([5.0])), n)),
             Page(keys, RunLengthEncodedBlock(LongArrayBlock(DOUBLE, np.array([6.0])), n))]
    plan = plan_text([BIGINT, DOUBLE], step="single", groupByChannels=[0],
                     aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT)], expectedGroups=16)
    act, st = run_gpu_stats(plan, pages)
    assert_rows_equal_ignore_order(act, run_oracle(plan, pages), nkeys=1)
    assert st.super_pages == 4                      # [d1, d1again] [d2] [rle 5.0] [rle 6.0]


def test_varchar_keys_across_pages():
    # var-width blocks of different byte lengths: offsets are rebased onto the super-page, payloads concatenated
    V8 = createVarcharType(8)
    rng = np.random.default_rng(5)
    words = ["", "a", "bb", "ccc", "dddd", "eeeee", "ffffff", "ggggggg", "hhhhhhhh", None]
    pages = []
    for n in (1, 777, 4096, 13, 2500):
        ks = [words[int(i)] for i in rng.integers(0, len(words), n)]
        pages.append(Page(VariableWidthBlock.from_values(V8, ks), LongArrayBlock(BIGINT, rng.integers(-1000, 1000, n))))
    # a block that is a slice of a larger one: its first offset is not 0
    big = VariableWidthBlock.from_values(V8, ["zz", "yyy", "a", "bb", "hhhhhhhh"])
    sl = VariableWidthBlock(V8, big.data, big.offsets[2:].copy())
    pages.append(Page(sl, LongArrayBlock(BIGINT, [1, 2, 3])))
    plan = plan_text([V8, BIGINT], step="single", groupByChannels=[0],
                     aggregations=[Aggregation("sum", BIGINT, 1, BIGINT), Aggregation("count", BIGINT)], expectedGroups=64)
    act, st = run_gpu_stats(plan, pages)
    assert_rows_equal_ignore_order(act, run_oracle(plan, pages), nkeys=1)
    assert st.super_pages <= 3                      # pages with and without a null vector are different signatures


@pytest.mark.parametrize("page_fn,plan_fn,nkeys", [(tpch.q1_page, tpch.q1_plan, 2), (tpch.q1_page_encoded, tpch.q1_plan, 2),
                                                   (tpch.highcard_page, lambda: tpch.highcard_plan(1000), 1)])
def test_serialized_small_pages_are_coalesced(page_fn, plan_fn, nkeys):
    total = 120_011
    pages = cut(page_fn, total, random_sizes(total, 1, 8192, 3))
    plan = plan_fn()
    act, st = run_gpu_stats(plan, pages, serialized=True)
    assert_rows_equal_ignore_order(act, run_oracle(plan, pages), nkeys=nkeys, rtol=sequential_bound(total))
    assert st.super_pages == 1 and st.input_pages == len(pages)


def test_serialized_pages_with_and_without_nulls_interleave():
    # blocks with null flags take the device decode path, the others the appender: both orders, same table
    rng = np.random.default_rng(9)
    pages = []
    for i in range(12):
        n = int(rng.integers(1, 3000))
        k = rng.integers(0, 50, n)
        v = rng.random(n)
        nulls = (rng.random(n) < 0.2).astype(np.uint8) if i % 3 == 1 else None
        pages.append(Page(LongArrayBlock(BIGINT, k), LongArrayBlock(DOUBLE, v, nulls)))
    plan = plan_text([BIGINT, DOUBLE], step="single", groupByChannels=[0],
                     aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT, 1, DOUBLE)], expectedGroups=64)
    act, st = run_gpu_stats(plan, pages, serialized=True)
    assert_rows_equal_ignore_order(act, run_oracle(plan, pages), nkeys=1)


def test_partial_flush_counts_rows_waiting_in_the_slot():
    # PARTIAL step with a small memory limit: the operator must report "full" although the rows are still in the staging slot
    total = 60_000
    pages = cut(tpch.highcard_page, total, [3000] * 20)
    plan = plan_text(tpch.HC_INPUT, step="partial", groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)],
                     expectedGroups=1000, maxPartialGroups=2000)
    op = GpuPipelineOperator(plan, 0)
    try:
        out = toPages(op, pages)
        assert len(out) >= 5                        # several partial flushes
    finally:
        op.close()
    fplan = plan_text([BIGINT, BIGINT, DOUBLE], step="final", groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)],
                      expectedGroups=20_000)
    splan = tpch.highcard_plan(20_000)
    from util import run_gpu
    assert_rows_equal_ignore_order(run_gpu(fplan, out), run_oracle(splan, pages), nkeys=1)


def test_reset_drops_rows_waiting_in_the_slot():
    total = 50_000
    pages = cut(tpch.q6_page, total, [10_000] * 5)
    op = GpuPipelineOperator(tpch.q6_plan(), 0)
    try:
        op.addInput(pages[0])
        op.addInput(pages[1])
        op.reset()                                  # nothing of the first two pages may survive
        act = rows_of(toPages(op, pages[2:]))
        exact, _ = exact_q6(pages[2:])
        assert_rows_equal_ignore_order(act, exact)
    finally:
        op.close()
