"""The page wire format (PagesSerdeUtil + the spi block encodings): the oracle's writer/reader against the reference's
serialized-size known answers (main/test/execution/buffer/TestPagesSerde.java:64-108) and round trips; the GPU ingest of
serialized pages (pgpu_add_serialized_page) against the same pages fed block by block."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import pages_serde as S  # noqa: E402

from presto_b200 import tpch  # noqa: E402
from presto_b200.spi import (BIGINT, BOOLEAN, DATE, DOUBLE, INTEGER, VARCHAR, ByteArrayBlock, DictionaryBlock, Int128ArrayBlock, IntArrayBlock,  # noqa: E402
                             LongArrayBlock, Page, RunLengthEncodedBlock, VariableWidthBlock, block_of, createDecimalType, createVarcharType)

FRAME = 13  # positionCount, markers, uncompressedSize, size (PagesSerdeUtil.writeSerializedPage)


def test_bigint_serialized_size_kat():
    # TestPagesSerde.testBigintSerializedSize: an empty builder builds RLE(null) x 0 -> 52 bytes; values: 35 + 9, then + 8
    empty = RunLengthEncodedBlock(LongArrayBlock(BIGINT, [0], [1]), 0)
    assert len(S.write_serialized_page([empty], 0)) == 52
    assert len(S.write_serialized_page([LongArrayBlock(BIGINT, [123])], 1)) == 35 + 9
    assert len(S.write_serialized_page([LongArrayBlock(BIGINT, [123, 456])], 2)) == 35 + 9 + 8


def test_varchar_serialized_size_kat():
    # TestPagesSerde.testVarcharSerializedSize: 44 empty, + 4 + 5 for "alice", + 4 + 3 for "bob"
    assert len(S.write_serialized_page([VariableWidthBlock.from_values(VARCHAR, [])], 0)) == 44
    assert len(S.write_serialized_page([VariableWidthBlock.from_values(VARCHAR, ["alice"])], 1)) == 44 + 9
    assert len(S.write_serialized_page([VariableWidthBlock.from_values(VARCHAR, ["alice", "bob"])], 2)) == 44 + 9 + 7


def test_null_bits_are_msb_first():
    # EncoderUtil.encodeNullsAsBits: position 0 -> 0b1000_0000; the ragged tail fills the high bits of the last byte
    assert S.encode_nulls_as_bits(np.array([1, 0, 0, 0, 0, 0, 0, 1, 0, 1], dtype=np.uint8), 10) == bytes([1, 0b10000001, 0b01000000])
    assert S.encode_nulls_as_bits(None, 10) == b"\x00"


def sample_blocks(n, seed=7):
    rng = np.random.default_rng(seed)
    nulls = (rng.random(n) < 0.3).astype(np.uint8)
    words = ["", "a", "bc", "def", "ghij"]
    return [
        LongArrayBlock(BIGINT, rng.integers(-2**62, 2**62, n)),
        LongArrayBlock(BIGINT, rng.integers(-2**62, 2**62, n), nulls),
        LongArrayBlock(DOUBLE, rng.random(n) * 100, nulls),
        IntArrayBlock(DATE, rng.integers(0, 20000, n)),
        IntArrayBlock(INTEGER, rng.integers(-2**31, 2**31 - 1, n), nulls),
        ByteArrayBlock(BOOLEAN, rng.integers(0, 2, n), nulls),
        Int128ArrayBlock(createDecimalType(30, 2), [int(x) * 10**12 for x in rng.integers(-10**9, 10**9, n)], nulls),
        VariableWidthBlock.from_values(createVarcharType(4), [None if nulls[i] else words[int(x)] for i, x in enumerate(rng.integers(0, 5, n))]),
        VariableWidthBlock.from_values(createVarcharType(4), [words[int(x)] for x in rng.integers(0, 5, n)]),
        DictionaryBlock(LongArrayBlock(DOUBLE, [0.0, 0.01, 0.02, 0.03]), rng.integers(0, 4, n)),
        DictionaryBlock(VariableWidthBlock.from_values(createVarcharType(1), ["A", "N", None]), rng.integers(0, 3, n)),
        RunLengthEncodedBlock(LongArrayBlock(BIGINT, [42]), n),
    ]


def storage_values(block):
    """what the wire carries: raw storage words / bytes per position"""
    if isinstance(block, DictionaryBlock):
        d = storage_values(block.dictionary)
        return [d[i] for i in block.ids.tolist()]
    if isinstance(block, RunLengthEncodedBlock):
        return storage_values(block.value) * block.position_count
    n = block.position_count
    nulls = block.valueIsNull
    if isinstance(block, VariableWidthBlock):
        raw = block.data.tobytes()
        vals = [raw[block.offsets[i]:block.offsets[i + 1]] for i in range(n)]
    elif isinstance(block, Int128ArrayBlock):
        vals = [bytes(r) for r in block.values.view(np.uint8).reshape(n, 16)]
    else:
        vals = block.values.tolist()
    return [None if nulls is not None and nulls[i] else vals[i] for i in range(n)]


@pytest.mark.parametrize("n", [0, 1, 7, 8, 9, 64, 1000])
def test_round_trip_every_encoding(n):
    blocks = sample_blocks(n)
    data = S.write_raw_page(blocks)
    back = S.read_raw_page(data)
    assert len(back) == len(blocks)
    for b, r in zip(blocks, back):
        assert r["positionCount"] == n
        assert S.block_values(r) == storage_values(b)


def test_non_null_values_only_are_written():
    # LongArrayBlockEncoding.writeBlock: a block that may have nulls carries nonNullCount + the non-null values
    b = LongArrayBlock(BIGINT, [1, 2, 3, 4], [0, 1, 1, 0])
    data = S.write_block(b)
    assert len(data) == 4 + 10 + 4 + 1 + 1 + 4 + 2 * 8
    assert data[-16:] == np.array([1, 4], dtype="<i8").tobytes()


# ---------------------------------------------------------------------------------------------------- GPU ingest
def _drain(op):
    from util import rows_of
    out = []
    while not op.isFinished():
        p = op.getOutput()
        if p is not None:
            out.append(p)
    return rows_of(out)


def _run_both(plan, pages):
    """the same pages through pgpu_add_page (blocks) and pgpu_add_serialized_page (wire format)"""
    from presto_b200.operators import GpuPipelineOperator
    results = []
    for serialized in (False, True):
        op = GpuPipelineOperator(plan, 0)
        try:
            for pg in pages:
                if serialized:
                    op.addSerializedInput(S.write_raw_page(pg.blocks), pg.position_count)
                else:
                    op.addInput(pg)
            op.finish()
            results.append(_drain(op))
        finally:
            op.close()
    return results


@pytest.mark.gpu
def test_serialized_q1_pages_match_block_pages():
    from util import assert_rows_equal_ignore_order, run_oracle, sequential_bound
    total = 300_000
    pages = [tpch.q1_page(0, 100_000, total), tpch.q1_page(100_000, 200_000, total)]
    blocks, wire = _run_both(tpch.q1_plan(), pages)
    assert_rows_equal_ignore_order(wire, blocks, nkeys=2, rtol=1e-12)
    assert_rows_equal_ignore_order(wire, run_oracle(tpch.q1_plan(), pages), nkeys=2, rtol=sequential_bound(total))


@pytest.mark.gpu
def test_serialized_dictionary_and_rle_pages():
    from util import assert_rows_equal_ignore_order
    total = 200_000
    pages = [tpch.q1_page_encoded(0, total, total)]   # DictionaryBlocks + a RunLengthEncodedBlock
    blocks, wire = _run_both(tpch.q1_plan(), pages)
    assert_rows_equal_ignore_order(wire, blocks, nkeys=2, rtol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 7, 8, 9, 1023, 1024, 1025, 70_001])
def test_serialized_pages_with_nulls(n):
    """null bit streams and non-null-only value arrays are re-expanded on the device: GROUP BY a nullable VARCHAR key,
    aggregates over nullable BIGINT / DOUBLE / INTEGER / long DECIMAL channels"""
    from presto_b200.operators import Aggregation, plan_text
    from util import assert_rows_equal_ignore_order, run_oracle
    rng = np.random.default_rng(n)
    nulls = lambda frac: (rng.random(n) < frac).astype(np.uint8)  # noqa: E731
    words = ["a", "bc", "def", ""]
    knull = nulls(0.2)
    dec = createDecimalType(30, 2)
    key = VariableWidthBlock.from_values(createVarcharType(4), [None if knull[i] else words[int(x)] for i, x in enumerate(rng.integers(0, 4, n))])
    page = Page(key,
                LongArrayBlock(BIGINT, rng.integers(-10**6, 10**6, n), nulls(0.3)),
                LongArrayBlock(DOUBLE, rng.integers(0, 1000, n).astype(np.float64), nulls(0.5)),
                IntArrayBlock(INTEGER, rng.integers(-1000, 1000, n), nulls(0.9)),
                Int128ArrayBlock(dec, [int(x) for x in rng.integers(-10**9, 10**9, n)], nulls(0.1)),
                ByteArrayBlock(BOOLEAN, rng.integers(0, 2, n), nulls(0.4)))
    types = [createVarcharType(4), BIGINT, DOUBLE, INTEGER, dec, BOOLEAN]
    plan = plan_text(types, step="single", groupByChannels=[0], aggregations=[
        Aggregation("sum", BIGINT, 1, BIGINT), Aggregation("count", BIGINT, 1, BIGINT), Aggregation("sum", DOUBLE, 2, DOUBLE),
        Aggregation("max", INTEGER, 3, INTEGER), Aggregation("sum", createDecimalType(38, 2), 4, dec), Aggregation("count", BIGINT, 5, BOOLEAN),
        Aggregation("count", BIGINT)], expectedGroups=16)
    blocks, wire = _run_both(plan, [page])
    assert_rows_equal_ignore_order(wire, blocks, nkeys=1)
    assert_rows_equal_ignore_order(wire, run_oracle(plan, [page]), nkeys=1)


@pytest.mark.gpu
def test_serialized_page_errors():
    from presto_b200.operators import GpuPipelineOperator, TrinoException
    op = GpuPipelineOperator(tpch.q6_plan(), 0)
    try:
        page = tpch.q6_page(0, 1000, 1000)
        good = S.write_raw_page(page.blocks)
        with pytest.raises(TrinoException):
            op.addSerializedInput(good[:-5], 1000)              # truncated
        with pytest.raises(TrinoException):
            op.addSerializedInput(good, 999)                    # position count differs
        with pytest.raises(TrinoException):
            op.addSerializedInput(S.write_raw_page(page.blocks[:-1]), 1000)   # channel count differs
        with pytest.raises(TrinoException):
            op.addSerializedInput(S.write_raw_page(list(page.blocks[1:]) + [page.blocks[0]]), 1000)  # encodings do not match the types
        op.addSerializedInput(good, 1000)                       # the operator is still usable
        op.finish()
        assert len(_drain(op)) == 1
    finally:
        op.close()


@pytest.mark.gpu
def test_corrupt_serialized_blocks_are_errors_not_faults():
    """header fields the device would follow blindly are checked on the host first: the non-null count against the null bits, the
    end offsets of VARIABLE_WIDTH blocks, dictionary ids against the dictionary (the reference fails with index errors in readBlock)"""
    import struct
    from presto_b200.operators import Aggregation, GpuPipelineOperator, TrinoException, plan_text
    n = 1000
    rng = np.random.default_rng(1)
    nulls = (rng.random(n) < 0.3).astype(np.uint8)
    v8 = createVarcharType(8)
    words = ["", "a", "bcd", "efghijkl"]
    key = VariableWidthBlock.from_values(v8, [words[int(x)] for x in rng.integers(0, 4, n)])
    dict_block = DictionaryBlock(LongArrayBlock(BIGINT, np.arange(10)), rng.integers(0, 10, n).astype(np.int32))
    page = Page(key, LongArrayBlock(BIGINT, rng.integers(0, 100, n), nulls), dict_block)
    plan = plan_text([v8, BIGINT, BIGINT], step="single", groupByChannels=[0], aggregations=[Aggregation("sum", BIGINT, 1, BIGINT), Aggregation("sum", BIGINT, 2, BIGINT)], expectedGroups=16)
    good = bytearray(S.write_raw_page(page.blocks))

    def at(name, nth=0):
        i = -1
        for _ in range(nth + 1):
            i = good.index(name, i + 1)
        return i + len(name)

    cases = {}
    # VARIABLE_WIDTH: [positionCount][n end offsets]...: a descending offset, and one beyond the data
    vw = at(b"VARIABLE_WIDTH")
    bad = bytearray(good); struct.pack_into("<i", bad, vw + 4 + 4 * 10, 0); cases["offsets descend"] = bad
    bad = bytearray(good); struct.pack_into("<i", bad, vw + 4 + 4 * (n - 1), 1 << 30); cases["offset beyond the data"] = bad
    # LONG_ARRAY with nulls: [positionCount][hasNulls = 1][bits][nonNull count]...: flip a null bit / change the count
    la = at(b"LONG_ARRAY")
    bad = bytearray(good); bad[la + 4 + 1 + 3] ^= 0x10; cases["null bit flipped"] = bad
    cnt = la + 4 + 1 + (n + 7) // 8
    bad = bytearray(good); struct.pack_into("<i", bad, cnt, struct.unpack_from("<i", good, cnt)[0] - 1); cases["non-null count lowered"] = bad
    # DICTIONARY: [positionCount][dictionary block][n ids][24 bytes id]: an id outside the dictionary, and a negative one
    ids = len(good) - 24 - 4 * n
    bad = bytearray(good); struct.pack_into("<i", bad, ids + 4 * 7, 10); cases["id == dictionary size"] = bad
    bad = bytearray(good); struct.pack_into("<i", bad, ids + 4 * 7, -1); cases["negative id"] = bad
    op = GpuPipelineOperator(plan, 0)
    try:
        for name, payload in cases.items():
            with pytest.raises(TrinoException) as e:
                op.addSerializedInput(bytes(payload), n)
            assert "corrupt" in str(e.value), (name, str(e.value))
        op.addSerializedInput(bytes(good), n)     # the pipeline is still usable and the device never saw the bad pages
        op.finish()
        from util import assert_rows_equal_ignore_order, run_oracle
        assert_rows_equal_ignore_order(_drain(op), run_oracle(plan, [page]), nkeys=1)
    finally:
        op.close()


# ------------------------------------------------------------------------------------------- GPU output in wire format
def _decoded_rows(blocks):
    cols = [S.block_values(b) for b in blocks]
    return list(zip(*cols)) if cols else []


def _drain_serialized(op):
    out = []
    while not op.isFinished():
        r = op.getSerializedOutput()
        if r is not None:
            payload, n = r
            blocks = S.read_raw_page(payload)
            assert all(b["positionCount"] == n for b in blocks)
            out.append((payload, n, blocks))
    return out


def _nullable_page(n, seed):
    rng = np.random.default_rng(seed)
    nulls = lambda frac: (rng.random(n) < frac).astype(np.uint8)  # noqa: E731
    words = ["a", "bc", "def", ""]
    knull = nulls(0.2)
    dec = createDecimalType(30, 2)
    key = VariableWidthBlock.from_values(createVarcharType(4), [None if knull[i] else words[int(x)] for i, x in enumerate(rng.integers(0, 4, n))])
    return Page(key,
                LongArrayBlock(BIGINT, rng.integers(0, 50, n), nulls(0.1)),
                LongArrayBlock(DOUBLE, rng.integers(0, 1000, n).astype(np.float64), nulls(0.5)),
                IntArrayBlock(INTEGER, rng.integers(-1000, 1000, n), nulls(0.97)),
                Int128ArrayBlock(dec, [int(x) for x in rng.integers(-10**9, 10**9, n)], nulls(0.1)),
                ByteArrayBlock(BOOLEAN, rng.integers(0, 2, n), nulls(0.4)))


@pytest.mark.gpu
@pytest.mark.parametrize("n,max_rows", [(1, 0), (9, 0), (5000, 0), (5000, 37), (70_001, 2048)])
@pytest.mark.parametrize("step", ["single", "partial"])
def test_serialized_output_equals_block_output(n, max_rows, step):
    """GROUP BY (nullable varchar, nullable bigint): var-width and fixed-width keys with nulls, nullable aggregate results of every
    width (1 / 4 / 8 / 16 bytes), PARTIAL states (incl. the decimal VARBINARY state), several output pages"""
    from presto_b200.operators import Aggregation, GpuPipelineOperator, plan_text
    dec = createDecimalType(30, 2)
    types = [createVarcharType(4), BIGINT, DOUBLE, INTEGER, dec, BOOLEAN]
    plan = plan_text(types, step=step, groupByChannels=[0, 1], aggregations=[
        Aggregation("sum", DOUBLE, 2, DOUBLE), Aggregation("max", INTEGER, 3, INTEGER), Aggregation("sum", createDecimalType(38, 2), 4, dec),
        Aggregation("min", BOOLEAN, 5, BOOLEAN), Aggregation("avg", DOUBLE, 2, DOUBLE), Aggregation("count", BIGINT)], expectedGroups=1024, maxOutputRows=max_rows)
    page = _nullable_page(n, n)
    want, got = [], []
    for wire in (False, True):
        op = GpuPipelineOperator(plan, 0)
        try:
            op.addInput(page)
            op.finish()
            if wire:
                pages = _drain_serialized(op)
                if max_rows:   # at most 5 x 51 = 255 groups: (5000, 37) splits into several pages, (70_001, 2048) stays one
                    total = sum(k for _, k, _ in pages)
                    assert all(k <= max_rows for _, k, _ in pages) and len(pages) == -(-total // max_rows)
                for _, _, blocks in pages:
                    got += _decoded_rows(blocks)
            else:
                while not op.isFinished():
                    p = op.getOutput()
                    if p is not None:
                        want += _decoded_rows(S.read_raw_page(S.write_raw_page(p.blocks)))   # same storage view of the block output
        finally:
            op.close()
    assert len(got) == len(want) > 0
    assert sorted(map(repr, got)) == sorted(map(repr, want))


@pytest.mark.gpu
def test_partial_output_travels_in_wire_format_to_final():
    """PARTIAL on the GPU -> serialized pages -> FINAL on the GPU (pgpu_add_serialized_page) == SINGLE step"""
    from presto_b200.operators import GpuPipelineOperator
    from test_oracle_protocol import q1_final_plan
    from util import assert_rows_equal_ignore_order
    total = 400_000
    pages = [tpch.q1_page(i * 100_000, 100_000, total) for i in range(4)]
    final = GpuPipelineOperator(q1_final_plan(), 0)
    try:
        for half in (pages[:2], pages[2:]):           # two "workers"
            partial = GpuPipelineOperator(tpch.q1_plan("partial"), 0)
            try:
                for pg in half:
                    partial.addInput(pg)
                partial.finish()
                for payload, n, _ in _drain_serialized(partial):
                    final.addSerializedInput(payload, n)
            finally:
                partial.close()
        final.finish()
        rows = _drain(final)
    finally:
        final.close()
    single = GpuPipelineOperator(tpch.q1_plan(), 0)
    try:
        for pg in pages:
            single.addInput(pg)
        single.finish()
        assert_rows_equal_ignore_order(rows, _drain(single), nkeys=2, rtol=1e-12)
    finally:
        single.close()


@pytest.mark.gpu
def test_serialized_output_of_filter_and_project():
    """the filter + project-only mode (no aggregation): every input page yields one compacted output page"""
    from presto_b200.operators import GpuPipelineOperator, plan_text
    from presto_b200.relational import call, compare, constant, field
    plan = plan_text([BIGINT, DOUBLE], compare("lt", field(0, BIGINT), constant(500, BIGINT)),
                     [field(0, BIGINT), call("multiply", DOUBLE, field(1, DOUBLE), constant(2.0, DOUBLE))])
    rng = np.random.default_rng(5)
    n = 10_000
    page = Page(LongArrayBlock(BIGINT, rng.integers(0, 1000, n), (rng.random(n) < 0.2).astype(np.uint8)),
                LongArrayBlock(DOUBLE, rng.random(n), (rng.random(n) < 0.3).astype(np.uint8)))
    outs = []
    for wire in (False, True):
        op = GpuPipelineOperator(plan, 0)
        try:
            op.addInput(page)
            if wire:
                payload, k = op.getSerializedOutput()
                outs.append(_decoded_rows(S.read_raw_page(payload)))
            else:
                outs.append(_decoded_rows(S.read_raw_page(S.write_raw_page(op.getOutput().blocks))))
        finally:
            op.close()
    assert outs[0] == outs[1] and 0 < len(outs[0]) < n      # order preserving compaction: same rows in the same order


@pytest.mark.gpu
def test_compressed_serialized_pages():
    """SerializedPages with the COMPRESSED marker (PagesSerde.serialize :96-118: the payload as one LZ4 block, kept only when it
    shrinks below 80 %): inflated on the host as PagesSerde.deserialize :156-165 does. Compressed here with Arrow's LZ4 codec, an
    implementation independent of the library's decoder."""
    pa = pytest.importorskip("pyarrow")
    from presto_b200.operators import GpuPipelineOperator, TrinoException
    from util import assert_rows_equal_ignore_order
    total = 250_000
    pages = [tpch.q1_page(0, 100_000, total), tpch.q1_page_encoded(100_000, 150_000, total)]
    blocks, _ = _run_both(tpch.q1_plan(), pages)
    codec = pa.Codec("lz4_raw")
    op = GpuPipelineOperator(tpch.q1_plan(), 0)
    try:
        shrunk = 0
        for pg in pages:
            raw = S.write_raw_page(pg.blocks)
            packed = codec.compress(raw, asbytes=True)
            if len(packed) <= 0.8 * len(raw):                 # PagesSerde.MINIMUM_COMPRESSION_RATIO
                op.addSerializedPage(packed, pg.position_count, markers=1, uncompressedSizeInBytes=len(raw))
                shrunk += 1
            else:
                op.addSerializedPage(raw, pg.position_count, markers=0, uncompressedSizeInBytes=len(raw))
        assert shrunk >= 1
        op.finish()
        got = _drain(op)
    finally:
        op.close()
    assert_rows_equal_ignore_order(got, blocks, nkeys=2, rtol=1e-12)
    # wrong sizes, damaged slices and encrypted pages are errors
    op = GpuPipelineOperator(tpch.q1_plan(), 0)
    try:
        raw = S.write_raw_page(pages[0].blocks)
        packed = codec.compress(raw, asbytes=True)
        for slice_, markers, size in ((packed, 1, len(raw) - 1), (packed[:len(packed) // 2], 1, len(raw)), (raw, 1, len(raw)), (raw, 2, len(raw)), (raw, 8, len(raw))):
            with pytest.raises(TrinoException):
                op.addSerializedPage(slice_, pages[0].position_count, markers=markers, uncompressedSizeInBytes=size)
        assert op.needsInput()
    finally:
        op.close()


def test_lz4_slices_are_decoded_on_the_host_without_a_gpu():
    """a compile-only pipeline (device -1) gets as far as the host-side LZ4 decode: a good slice then fails with "compile-only",
    a damaged one with "corrupt page" — pins the block decoder against Arrow's encoder on the CPU"""
    pa = pytest.importorskip("pyarrow")
    from presto_b200.operators import GpuPipelineOperator, TrinoException
    rng = np.random.default_rng(3)
    page = tpch.q1_page(0, 50_000, 50_000)
    raw = S.write_raw_page(page.blocks)
    packed = pa.Codec("lz4_raw").compress(raw, asbytes=True)
    assert len(packed) < len(raw)
    op = GpuPipelineOperator(tpch.q1_plan(), -1)
    try:
        with pytest.raises(TrinoException) as e:
            op.addSerializedPage(packed, page.position_count, markers=1, uncompressedSizeInBytes=len(raw))
        assert "compile-only" in str(e.value)
        for trial in range(20):
            bad = bytearray(packed)
            for at in rng.integers(0, len(bad), 4):
                bad[at] ^= int(rng.integers(1, 256))
            with pytest.raises(TrinoException) as e:
                op.addSerializedPage(bytes(bad), page.position_count, markers=1, uncompressedSizeInBytes=len(raw))
            assert "corrupt page" in str(e.value) or "compile-only" in str(e.value)   # (a flipped literal byte still decodes)
        with pytest.raises(TrinoException) as e:
            op.addSerializedPage(packed, page.position_count, markers=2, uncompressedSizeInBytes=len(raw))
        assert "encrypted" in str(e.value)
    finally:
        op.close()
