"""Parquet column chunks through pgpu_add_parquet_row_group on the GPU against (a) what pyarrow's reader returns for the same
files (tests/golden/*.expected.npz), (b) the CPU oracle operator fed with plain pages of those values, and (c) the GPU path
itself fed with the same plain pages. Bit-exact: identity projections return the file's values; aggregates over exact data."""
import io
import os
import sys

import numpy as np
import pytest

from parquet_util import GOLDEN, chunks_of, dictionary_chunk, expected_of
from presto_b200 import tpch
from presto_b200.operators import Aggregation, GpuPipelineOperator, TrinoException, plan_text
from presto_b200.relational import compare, constant, field
from presto_b200.spi import (BIGINT, BOOLEAN, DATE, DOUBLE, INTEGER, ByteArrayBlock, IntArrayBlock, LongArrayBlock, Page, VariableWidthBlock,
                             createVarcharType)
from util import assert_rows_equal_ignore_order, assert_rows_equal_ordered, rows_of, run_gpu, run_oracle, sequential_bound

pytestmark = pytest.mark.gpu
V1, V3 = createVarcharType(1), createVarcharType(3)
NULLABLE = [("k", BIGINT), ("runs", INTEGER), ("x", DOUBLE), ("flag", V3), ("b", BOOLEAN), ("allnull", BIGINT), ("big", BIGINT)]


def run_parquet(plan, groups, columns, device=0):
    """groups = chunks_of(file); columns = parquet column names in plan channel order"""
    op = GpuPipelineOperator(plan, device)
    out = []
    try:
        for rows, cols in groups:
            assert op.needsInput()
            op.addParquetRowGroup([cols[c] for c in columns], rows)
            p = op.getOutput()
            if p is not None and p.position_count:
                out.append(p)
        op.finish()
        while not op.isFinished():
            p = op.getOutput()
            if p is not None and p.position_count:
                out.append(p)
        return rows_of(out), op.stats()
    finally:
        op.close()


def plain_block(t, values, nulls):
    nn = nulls.astype(np.uint8) if nulls.any() else None
    if t.is_var_width:
        return VariableWidthBlock.from_values(t, [None if z else str(v) for v, z in zip(values, nulls)])
    if t.kind == 0:
        return ByteArrayBlock(t, values.astype(np.int8), nn)
    if t.kind in (2, 3):
        return IntArrayBlock(t, values.astype(np.int32), nn)
    if t.kind == 4:
        return LongArrayBlock(t, values.astype(np.float64), nn)
    return LongArrayBlock(t, values.astype(np.int64), nn)


def plain_pages(exp, schema, row_counts):
    pages, at = [], 0
    for n in row_counts:
        pages.append(Page(*[plain_block(t, exp[c + ".values"][at:at + n], exp[c + ".nulls"][at:at + n]) for c, t in schema]))
        at += n
    return pages


def test_identity_projection_returns_the_files_values():
    """every decode path at once: required / optional columns, dictionary ids with a NULL entry, RLE + bit-packed runs, PLAIN
    doubles with NULLs, RLE booleans, an all-NULL column, data page v2"""
    groups = chunks_of(os.path.join(GOLDEN, "nullable_v2.parquet"))
    exp = expected_of("nullable_v2")
    types = [t for _, t in NULLABLE]
    plan = plan_text(types, None, [field(i, t) for i, t in enumerate(types)])
    got, st = run_parquet(plan, groups, [c for c, _ in NULLABLE])
    want = rows_of(plain_pages(exp, NULLABLE, [g[0] for g in groups]))
    assert len(want) == 1000
    assert_rows_equal_ordered(got, want)
    assert st.kernel_launches > len(NULLABLE)


def test_aggregation_over_nullable_chunks_equals_oracle():
    groups = chunks_of(os.path.join(GOLDEN, "nullable_v2.parquet"))
    exp = expected_of("nullable_v2")
    types = [t for _, t in NULLABLE]
    plan = plan_text(types, compare("ge", field(0, BIGINT), constant(-400, BIGINT)), [field(3, V3), field(1, INTEGER), field(2, DOUBLE), field(6, BIGINT), field(4, BOOLEAN)],
                     step="single", groupByChannels=[0, 1],
                     aggregations=[Aggregation("count", BIGINT), Aggregation("count", BIGINT, 2, DOUBLE), Aggregation("min", DOUBLE, 2, DOUBLE),
                                   Aggregation("max", BIGINT, 3, BIGINT), Aggregation("min", BOOLEAN, 4, BOOLEAN)], expectedGroups=64)
    got, _ = run_parquet(plan, groups, [c for c, _ in NULLABLE])
    pages = plain_pages(exp, NULLABLE, [g[0] for g in groups])
    assert_rows_equal_ignore_order(got, run_oracle(plan, pages), nkeys=2)
    assert_rows_equal_ignore_order(got, run_gpu(plan, pages), nkeys=2)


LINEITEM = {"returnflag": V1, "linestatus": V1, "quantity": DOUBLE, "extendedprice": DOUBLE, "discount": DOUBLE, "tax": DOUBLE, "shipdate": DATE,
            "orderkey": BIGINT}


@pytest.mark.parametrize("plan,columns,nkeys", [
    (tpch.q1_plan(), ["returnflag", "linestatus", "quantity", "extendedprice", "discount", "tax", "shipdate"], 2),
    (tpch.q6_plan(), ["extendedprice", "discount", "shipdate", "quantity"], None),
    (tpch.highcard_plan(1000), ["orderkey", "quantity"], 1)], ids=["q1", "q6", "highcard"])
def test_tpch_plans_over_lineitem_row_groups(plan, columns, nkeys):
    """Q1 / Q6 / GROUP BY orderkey straight off Parquet row groups (dictionary-encoded flag, discount, tax, quantity chunks reach
    the kernels as DictionaryBlocks) = the same plans over plain pages of the values pyarrow reads"""
    groups = chunks_of(os.path.join(GOLDEN, "lineitem_v1.parquet"))
    exp = expected_of("lineitem_v1")
    got, _ = run_parquet(plan, groups, columns)
    pages = plain_pages(exp, [(c, LINEITEM[c]) for c in columns], [g[0] for g in groups])
    rtol = sequential_bound(600)
    assert_rows_equal_ignore_order(got, run_oracle(plan, pages), nkeys=nkeys, rtol=rtol)
    assert_rows_equal_ignore_order(got, run_gpu(plan, pages), nkeys=nkeys, rtol=rtol)


def test_large_row_group_size_independent_properties():
    """1.2 M rows written by pyarrow (dictionary chunk that falls back to PLAIN, many pages): count, exact integer sums and the
    number of groups are known in closed form"""
    pa = pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    n = 1_200_000
    k = (np.arange(n, dtype=np.int64) * 7919) % 100_003
    v = (np.arange(n, dtype=np.int64) % 1000).astype(np.float64)
    t = pa.table({"k": pa.array(k, mask=(np.arange(n) % 97 == 0)), "v": pa.array(v)})
    buf = io.BytesIO()
    pq.write_table(t, buf, compression="NONE", use_dictionary=["k"], dictionary_pagesize_limit=1 << 16, row_group_size=500_000, data_page_version="2.0")
    groups = chunks_of(buf.getvalue())
    plan = plan_text([BIGINT, DOUBLE], step="single", groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT)],
                     expectedGroups=200_000)
    got, _ = run_parquet(plan, groups, ["k", "v"])
    mask = np.arange(n) % 97 == 0
    assert len(got) == len(np.unique(k[~mask])) + 1
    assert sum(r[2] for r in got) == n
    assert sum(r[1] for r in got) == float(v.sum())
    null_row = [r for r in got if r[0] is None]
    assert len(null_row) == 1 and null_row[0][2] == int(mask.sum()) and null_row[0][1] == float(v[mask].sum())


def test_rejected_and_corrupt_chunks_raise():
    groups = chunks_of(os.path.join(GOLDEN, "lineitem_v1.parquet"))
    rows, cols = groups[0]
    op = GpuPipelineOperator(tpch.q6_plan(), 0)
    try:
        with pytest.raises(TrinoException) as e:   # a DOUBLE channel fed with the INT64 chunk
            op.addParquetRowGroup([cols["orderkey"], cols["discount"], cols["shipdate"], cols["quantity"]], rows)
        assert e.value.code == "NOT_SUPPORTED"
        b, pt, md, n, tl = cols["extendedprice"]
        with pytest.raises(TrinoException) as e:
            op.addParquetRowGroup([(b[:len(b) // 2], pt, md, n), cols["discount"], cols["shipdate"], cols["quantity"]], rows)
        assert e.value.code != "NOT_SUPPORTED"
    finally:
        op.close()


def test_dictionary_id_out_of_range_is_caught_on_the_device():
    """a page that points past its dictionary (PrimitiveColumnReader would throw from Dictionary.decodeToLong): the kernel flags it,
    the operator reports it where device faults surface"""
    plan = plan_text([BIGINT], step="single", aggregations=[Aggregation("sum", BIGINT, 0, BIGINT), Aggregation("count", BIGINT)])
    good = dictionary_chunk([10, 20, 30], [0, 1, 2, 2, 1, 0, 0, 2, 1, 1, 2], 2)
    op = GpuPipelineOperator(plan, 0)
    try:
        op.addParquetRowGroup([(good, 2, 0, 11)], 11)
        op.finish()
        out = []
        while not op.isFinished():
            p = op.getOutput()
            if p is not None:
                out.append(p)
        assert rows_of(out) == [(10 * 3 + 20 * 4 + 30 * 4, 11)]
    finally:
        op.close()
    bad = dictionary_chunk([10, 20, 30], [0, 1, 2, 3, 1, 0, 0, 2, 1, 1, 2], 2)   # id 3 of a 3-entry dictionary
    op = GpuPipelineOperator(plan, 0)
    try:
        with pytest.raises(TrinoException) as e:
            op.addParquetRowGroup([(bad, 2, 0, 11)], 11)
            op.finish()
            while not op.isFinished():
                op.getOutput()
        assert "dictionary id" in str(e.value)
    finally:
        op.close()


@pytest.mark.parametrize("as_integer", [False, True], ids=["flba", "integers"])
def test_decimal_columns_identity(as_integer):
    """DECIMALs stored as FIXED_LEN_BYTE_ARRAY (short: 6 and 4 bytes -> int64; long: 11 bytes -> sign-magnitude int128) or as INT64 /
    INT32 (widened): the blocks must hold the unscaled values pyarrow's reader returns, dictionary-encoded and plain, with NULLs"""
    pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    import test_parquet_oracle as O
    from presto_b200.spi import createDecimalType
    t = O.decimal_table()
    buf = io.BytesIO()
    pq.write_table(t, buf, compression="NONE", use_dictionary=["discount"], data_page_size=2048, write_batch_size=64, store_decimal_as_integer=as_integer)
    back = pq.read_table(io.BytesIO(buf.getvalue()))
    names = ["price", "discount", "wide", "small"]
    types = [createDecimalType(12, 2), createDecimalType(12, 2), createDecimalType(25, 4), createDecimalType(7, 2)]
    plan = plan_text(types, None, [field(i, ty) for i, ty in enumerate(types)])
    got, _ = run_parquet(plan, chunks_of(buf.getvalue()), names)
    want = list(zip(*[O.unscaled(back.column(n)) for n in names]))
    assert_rows_equal_ordered(got, want)


def test_sf1_lineitem_parquet_file_equals_golden_answers():
    """the reference's own golden TPC-H answers (hive_tpch q01 / q06, computed on DECIMAL(12,2) columns) straight off a Parquet
    file of the sf1 table: FIXED_LEN_BYTE_ARRAY(6) decimals, dictionary-encoded flags / quantity / discount / tax, 1 M-row groups"""
    pytest.importorskip("pyarrow")
    import test_tpch_golden as G
    from parquet_util import Q1_COLUMNS, tpch_lineitem_parquet
    groups = chunks_of(tpch_lineitem_parquet(1.0, 1_000_000))
    assert sum(g[0] for g in groups) == 6_001_215
    got, st = run_parquet(tpch.q1_decimal_plan(), groups, Q1_COLUMNS)
    G.check_q1_decimal(got)
    got, _ = run_parquet(tpch.q6_decimal_plan(), groups, ["extendedprice", "discount", "shipdate", "quantity"])
    assert got == G.golden_rows("hive_tpch_q06.result", (4,))


@pytest.mark.parametrize("compression,codec", [("SNAPPY", 1), ("LZ4", 7), ("GZIP", 2), ("ZSTD", 6)])
def test_compressed_lineitem_parquet_file_equals_golden_answers(compression, codec):
    """the same golden answers off a SNAPPY- / LZ4_RAW-compressed file of the sf1 table: pages are
    inflated on the host (ParquetCompressionUtils.java:43-70), everything else is the uncompressed path"""
    pytest.importorskip("pyarrow")
    import test_tpch_golden as G
    from parquet_util import Q1_COLUMNS, tpch_lineitem_parquet
    data = tpch_lineitem_parquet(1.0, 1_000_000, compression=compression)
    groups = chunks_of(data)
    assert all(len(c) == 6 and c[5] == codec for _, cols in groups for c in cols.values())
    assert len(data) < 0.95 * len(tpch_lineitem_parquet(1.0, 1_000_000))   # it really is compressed (decimal digits do not pack far)
    got, st = run_parquet(tpch.q1_decimal_plan(), groups, Q1_COLUMNS)
    G.check_q1_decimal(got)
    got, _ = run_parquet(tpch.q6_decimal_plan(), groups, ["extendedprice", "discount", "shipdate", "quantity"])
    assert got == G.golden_rows("hive_tpch_q06.result", (4,))


def test_compressed_nullable_pages_equal_the_uncompressed_file():
    """NULLs, booleans, mixed dictionary / plain chunks, data pages v1 and v2, small pages: compressed files give the blocks the
    uncompressed file gives"""
    pytest.importorskip("pyarrow")
    import io
    import pyarrow.parquet as pq
    sys.path.insert(0, GOLDEN)
    import make_parquet_fixtures as M
    plan = plan_text([t for _, t in NULLABLE if t != V3], None, [field(i, t) for i, (_, t) in enumerate([c for c in NULLABLE if c[1] != V3])])
    names = [n for n, t in NULLABLE if t != V3]
    table = M.nullable_table(20_000, seed=9)
    for version in ("1.0", "2.0"):
        outs = []
        for compression in ("NONE", "SNAPPY", "LZ4", "GZIP", "ZSTD"):
            buf = io.BytesIO()
            pq.write_table(table, buf, compression=compression, data_page_version=version, data_page_size=3000, use_dictionary=["runs"], row_group_size=7000)
            got, _ = run_parquet(plan, chunks_of(buf.getvalue()), names)
            outs.append(got)
        assert all(o == outs[0] for o in outs[1:]) and len(outs[0]) == 20_000


def test_hadoop_framed_lz4_chunk_values():
    """codec LZ4 (5), Hadoop framing: the values that reach the kernels are the values that were written"""
    pytest.importorskip("pyarrow")
    from parquet_util import plain_int64_chunk
    rng = np.random.default_rng(6)
    values = np.repeat(rng.integers(-10**9, 10**9, 20_000), 7)
    chunk = plain_int64_chunk(values, True, pages=5)
    plan = plan_text([BIGINT], step="single", aggregations=[Aggregation("sum", BIGINT, 0, BIGINT), Aggregation("min", BIGINT, 0, BIGINT),
                                                            Aggregation("max", BIGINT, 0, BIGINT), Aggregation("count", BIGINT)])
    op = GpuPipelineOperator(plan, 0)
    try:
        op.addParquetRowGroup([(chunk, 2, 0, len(values), 0, 5)], len(values))
        op.finish()
        out = []
        while not op.isFinished():
            p = op.getOutput()
            if p is not None:
                out.append(p)
    finally:
        op.close()
    assert rows_of(out) == [(int(values.sum()), int(values.min()), int(values.max()), len(values))]

