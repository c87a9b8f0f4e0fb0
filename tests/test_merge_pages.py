"""MergePages behind the filter + project operator (operator/project/MergePages.java:87-164; the cases of
core/trino-main/src/test/java/io/trino/operator/project/TestMergePages.java) and pages cut at max-output-rows.
The CPU half pins the oracle's restatement to the reference's test cases; the GPU half compares the CUDA path with the
oracle page by page (same boundaries, same rows, bit-exact)."""
import numpy as np
import pytest

from presto_b200.operators import GpuPipelineOperator, plan_text, toPages
from presto_b200.relational import call, compare, constant, field
from presto_b200.spi import BIGINT, DOUBLE, INTEGER, IntArrayBlock, LongArrayBlock, Page

from oracle_op import OracleOperator

TYPES = [BIGINT, INTEGER, DOUBLE]
ROW_BYTES = (8 + 1) + (4 + 1) + (8 + 1)   # Block.getSizeInBytes per position: the value + one null byte per channel
MAX = 2**31 - 1


def sequence_page(n, start=0):
    """SequencePageBuilder.createSequencePage: channel c holds start + c*... a sequence per channel"""
    k = np.arange(start, start + n)
    return Page(LongArrayBlock(BIGINT, k), IntArrayBlock(INTEGER, (k + 100).astype(np.int32)), LongArrayBlock(DOUBLE, k.astype(np.float64) * 0.5))


def identity_plan(**opts):
    return plan_text(TYPES, None, [field(i, t) for i, t in enumerate(TYPES)], **opts)


def run_pages(op, pages):
    try:
        return toPages(op, pages)
    finally:
        op.close()


def shape(pages):
    return [p.position_count for p in pages]


def rows(pages):
    out = []
    for p in pages:
        out.extend(p.to_rows())
    return out


CASES = {
    # name: (options, input page sizes, expected output page sizes)
    "min_page_size_threshold": (dict(mergeMinBytes=10 * ROW_BYTES, mergeMinRows=MAX, mergeMaxBytes=MAX), [10], [10]),
    "min_row_count_threshold": (dict(mergeMinBytes=1 << 20, mergeMinRows=10, mergeMaxBytes=MAX), [10], [10]),
    "buffer_small_pages": (dict(mergeMinBytes=20 * ROW_BYTES + 1, mergeMinRows=21, mergeMaxBytes=MAX), [10, 10], [20]),
    "flush_on_big_page": (dict(mergeMinBytes=100 * ROW_BYTES, mergeMinRows=100, mergeMaxBytes=MAX), [10, 100], [10, 100]),
    "flush_on_full_page": (dict(mergeMinBytes=20 * ROW_BYTES, mergeMinRows=MAX, mergeMaxBytes=20 * ROW_BYTES), [10, 10, 10, 10], [20, 20]),
    "merging_off": (dict(), [3, 1, 7], [3, 1, 7]),
    # a page goes out as it is when it has min-rows rows OR min-bytes bytes (MergePages.java:134): min-bytes 0 passes everything
    "rows_only_passes_everything": (dict(mergeMinRows=50), [10, 10], [10, 10]),
    "tail_flushed_at_finish": (dict(mergeMinRows=50, mergeMinBytes=1 << 20), [10, 10, 60, 5, 5], [20, 60, 10]),
    "big_pages_pass_untouched": (dict(mergeMinRows=5, mergeMinBytes=1 << 20), [10, 10], [10, 10]),
    "default_session_values": (dict(mergeMinRows=256, mergeMinBytes=500_000), [100, 100, 56, 300, 1], [256, 300, 1]),
}


def input_pages(sizes):
    out, start = [], 0
    for n in sizes:
        out.append(sequence_page(n, start))
        start += n
    return out


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_merge_pages_reference_cases(name):
    opts, sizes, want = CASES[name]
    pages = input_pages(sizes)
    out = run_pages(OracleOperator(identity_plan(**opts)), pages)
    assert shape(out) == want
    assert rows(out) == rows(pages)


def test_oracle_max_output_rows_cuts_filter_project_pages():
    pages = input_pages([1000, 10])
    out = run_pages(OracleOperator(identity_plan(maxOutputRows=300)), pages)
    assert shape(out) == [300, 300, 300, 100, 10]
    assert rows(out) == rows(pages)


def test_merge_option_validation():
    for bad in (dict(mergeMinBytes=100, mergeMaxBytes=50),):
        with pytest.raises(Exception):
            OracleOperator(identity_plan(**bad))
    with pytest.raises(Exception):
        OracleOperator(plan_text([BIGINT], None, None, step="single", groupByChannels=[0], aggregations=[], mergeMinRows=5))


# ------------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_merge_pages_matches_oracle_page_by_page(name):
    opts, sizes, want = CASES[name]
    pages = input_pages(sizes)
    got = run_pages(GpuPipelineOperator(identity_plan(**opts), 0), pages)
    ref = run_pages(OracleOperator(identity_plan(**opts)), pages)
    assert shape(got) == shape(ref) == want
    for g, r in zip(got, ref):
        assert g.to_rows() == r.to_rows()


@pytest.mark.gpu
def test_gpu_merge_with_filter_and_growing_buffer():
    # a selective filter leaves a few rows of every page: the buffer grows across many pages (the arena is reallocated with
    # the buffered rows kept), a page that passes many rows flushes it, the tail goes out at finish
    rng = np.random.default_rng(7)
    pages, start = [], 0
    for i in range(60):
        n = int(rng.integers(1, 5000))
        pages.append(sequence_page(n, start))
        start += n
    pages.insert(30, sequence_page(40_000, start))
    filt = compare("eq", call("modulus", BIGINT, field(0, BIGINT), constant(97, BIGINT)), constant(3, BIGINT))
    plan = plan_text(TYPES, filt, [field(2, DOUBLE), field(0, BIGINT)], mergeMinRows=256, mergeMinBytes=500_000)
    op = GpuPipelineOperator(plan, 0)
    got = toPages(op, pages)
    merged = op.stats().pages_merged
    op.close()
    ref = run_pages(OracleOperator(plan), pages)
    assert shape(got) == shape(ref)
    assert len(got) < len(pages) // 4 and merged > 40
    for g, r in zip(got, ref):
        assert g.to_rows() == r.to_rows()


@pytest.mark.gpu
def test_gpu_max_output_rows_cuts_filter_project_pages():
    # a page that passes more rows than max-output-rows goes out in several getOutput calls; the operator refuses input in between
    pages = input_pages([1000, 10])
    plan = identity_plan(maxOutputRows=300)
    op = GpuPipelineOperator(plan, 0)
    op.addInput(pages[0])
    first = op.getOutput()
    assert first.position_count == 300 and not op.needsInput()
    with pytest.raises(Exception):
        op.addInput(pages[1])
    rest = []
    while not op.needsInput():
        rest.append(op.getOutput())
    assert shape(rest) == [300, 300, 100]
    op.addInput(pages[1])
    op.finish()
    tail = []
    while not op.isFinished():
        p = op.getOutput()
        if p is not None:
            tail.append(p)
    op.close()
    assert shape(tail) == [10]
    assert rows([first] + rest + tail) == rows(pages)
    ref = run_pages(OracleOperator(plan), pages)
    assert shape(ref) == [300, 300, 300, 100, 10]


@pytest.mark.gpu
def test_gpu_merge_and_max_output_rows_together():
    pages = input_pages([10] * 50 + [700])
    plan = identity_plan(mergeMinRows=400, mergeMinBytes=1 << 20, maxOutputRows=128)
    got = run_pages(GpuPipelineOperator(plan, 0), pages)
    ref = run_pages(OracleOperator(plan), pages)
    assert shape(got) == shape(ref) == [128, 128, 128, 116, 128, 128, 128, 128, 128, 60]
    assert rows(got) == rows(pages)


@pytest.mark.gpu
def test_gpu_merge_reset_drops_the_buffer():
    plan = identity_plan(mergeMinRows=100, mergeMinBytes=1 << 20)
    op = GpuPipelineOperator(plan, 0)
    op.addInput(sequence_page(10))
    assert op.getOutput() is None
    op.reset()
    op.addInput(sequence_page(5, 1000))
    assert op.getOutput() is None
    op.finish()
    out = []
    while not op.isFinished():
        p = op.getOutput()
        if p is not None:
            out.append(p)
    op.close()
    assert shape(out) == [5] and out[0].to_rows()[0][0] == 1000
