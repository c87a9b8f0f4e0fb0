"""Test helpers for Parquet column chunks: the chunk bytes and their metadata out of a file (footer read by pyarrow — the
reference reads it with MetadataReader.readFooter, lib/trino-parquet/.../reader/MetadataReader.java — the column chunk
itself is sliced out of the file untouched, as ParquetReader.readColumnChunk does, .../reader/ParquetReader.java:244-268)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def chunks_of(path):
    """-> [(row count, {column name: (chunk bytes, physical type, max definition level, num_values)})] per row group"""
    from presto_b200 import parquet
    return list(parquet.row_groups(path))


def expected_of(name):
    z = np.load(os.path.join(GOLDEN, name + ".expected.npz"), allow_pickle=False)
    return {k: z[k] for k in z.files}


# ---- hand-built chunks (Thrift compact page headers written field by field) ---------------------------------------------------
def _varint(v):
    out = bytearray()
    while True:
        if v < 0x80:
            out.append(v)
            return bytes(out)
        out.append((v & 0x7F) | 0x80)
        v >>= 7


def _zz(v):
    return _varint((v << 1) ^ (v >> 63))


def page_header(page_type, size, num_values, encoding, compressed_size=None):
    """parquet.thrift PageHeader for a dictionary page (type 2) or a v1 data page (type 0; RLE levels)"""
    h = b"\x15" + _zz(page_type) + b"\x15" + _zz(size) + b"\x15" + _zz(size if compressed_size is None else compressed_size)
    if page_type == 2:
        return h + b"\x4c" + b"\x15" + _zz(num_values) + b"\x15" + _zz(encoding) + b"\x00" + b"\x00"
    return h + b"\x2c" + b"\x15" + _zz(num_values) + b"\x15" + _zz(encoding) + b"\x15" + _zz(3) + b"\x15" + _zz(3) + b"\x00" + b"\x00"


def bit_packed(values, width):
    """one bit-packed hybrid run (header + groups of 8, least significant bit first)"""
    groups = (len(values) + 7) // 8
    bits = np.zeros(groups * 8 * width, np.uint8)
    for i, v in enumerate(values):
        for b in range(width):
            bits[i * width + b] = (v >> b) & 1
    return _varint((groups << 1) | 1) + np.packbits(bits, bitorder="little").tobytes()


def dictionary_chunk(dictionary, ids, width):
    """required INT64 column: PLAIN dictionary page + one RLE_DICTIONARY v1 data page whose ids are one bit-packed run"""
    d = np.asarray(dictionary, "<i8").tobytes()
    body = bytes([width]) + bit_packed(ids, width)
    return page_header(2, len(d), len(dictionary), 0) + d + page_header(0, len(body), len(ids), 8) + body


def lz4_hadoop_frame(raw, blocks=2, chunks_per_block=2):
    """Hadoop Lz4Codec framing, the deprecated Parquet codec LZ4 (5) as ParquetCompressionUtils.decompressFramed reads it:
    per block [uncompressed length, big endian] then per chunk [compressed length, big endian][raw LZ4 block]"""
    import struct
    import pyarrow as pa
    codec = pa.Codec("lz4_raw")
    out = bytearray()
    cut = [len(raw) * k // blocks for k in range(blocks + 1)]
    for a, b in zip(cut, cut[1:]):
        out += struct.pack(">i", b - a)
        sub = [a + (b - a) * k // chunks_per_block for k in range(chunks_per_block + 1)]
        for c, d in zip(sub, sub[1:]):
            packed = codec.compress(raw[c:d], asbytes=True)
            out += struct.pack(">i", len(packed)) + packed
    return bytes(out)


def plain_int64_chunk(values, lz4_hadoop=False, pages=3):
    """required INT64 column, PLAIN v1 data pages; with lz4_hadoop the page bodies are Hadoop-framed LZ4"""
    values = np.asarray(values, "<i8")
    out = b""
    cut = [len(values) * k // pages for k in range(pages + 1)]
    for a, b in zip(cut, cut[1:]):
        raw = values[a:b].tobytes()
        body = lz4_hadoop_frame(raw) if lz4_hadoop else raw
        out += page_header(0, len(raw), b - a, 0, len(body)) + body
    return out


# ---- the tpch connector's lineitem rows as a Parquet file (pyarrow's writer) ------------------------------------------------
def decimal_array(unscaled, precision, scale):
    """int64 unscaled values -> pyarrow decimal128(precision, scale) without a Python object per value"""
    import pyarrow as pa
    v = np.ascontiguousarray(unscaled, np.int64)
    words = np.empty((len(v), 2), np.int64)
    words[:, 0] = v
    words[:, 1] = v >> 63          # sign extension of the little-endian 128-bit value
    return pa.Array.from_buffers(pa.decimal128(precision, scale), len(v), [None, pa.py_buffer(words.tobytes())])


def tpch_lineitem_parquet(sf, row_group_rows, as_integer=False, first_order=0, n_orders=None, compression="NONE"):
    """Q1's columns of TPC-H lineitem (product-side host generator, pinned by the sf1 golden answers in test_tpch_golden.py) with
    DECIMAL(12,2) numerics, dictionary-encoded flags, DATE ship dates -> the bytes of an uncompressed Parquet file"""
    import io
    import pyarrow as pa
    import pyarrow.parquet as pq
    from presto_b200 import tpch
    if n_orders is None:
        n_orders = tpch.tpch_order_count(sf) - first_order
    c = lambda name, dec=False: tpch.tpch_host_column(name, sf, first_order, n_orders, dec)
    t = pa.table({
        "returnflag": pa.DictionaryArray.from_arrays(pa.array(c("returnflag_id").astype(np.int32)), pa.array(["A", "N", "R"])),
        "linestatus": pa.DictionaryArray.from_arrays(pa.array(c("linestatus_id").astype(np.int32)), pa.array(["F", "O"])),
        "quantity": decimal_array(c("quantity", True), 12, 2), "extendedprice": decimal_array(c("extendedprice", True), 12, 2),
        "discount": decimal_array(c("discount", True), 12, 2), "tax": decimal_array(c("tax", True), 12, 2),
        "shipdate": pa.array(c("shipdate").astype(np.int32), pa.date32())})
    buf = io.BytesIO()
    pq.write_table(t, buf, compression=compression, use_dictionary=["returnflag", "linestatus", "quantity", "discount", "tax"], row_group_size=row_group_rows,
                   write_statistics=False, store_decimal_as_integer=as_integer)
    return buf.getvalue()


Q1_COLUMNS = ["returnflag", "linestatus", "quantity", "extendedprice", "discount", "tax", "shipdate"]
