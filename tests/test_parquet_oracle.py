"""Pins the Parquet page oracle (oracle/parquet_pages.py) against an independent implementation: files written by pyarrow's
writer, expectations = what pyarrow's C++ reader returns (committed under tests/golden/ by make_parquet_fixtures.py, and
regenerated live with other writer settings when pyarrow is importable)."""
import io
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import parquet_pages as PP  # noqa: E402
from parquet_util import GOLDEN, chunks_of, dictionary_chunk, expected_of  # noqa: E402


def check_file(path_or_bytes, exp):
    row0 = 0
    seen = set()
    for rows, cols in chunks_of(path_or_bytes):
        for name, (buf, ptype, max_def, nvals, tl) in cols.items():
            assert nvals == rows
            values, nulls, info = PP.decode_chunk(buf, ptype, max_def, nvals, tl)
            e_nulls = exp[name + ".nulls"][row0:row0 + rows]
            e_vals = exp[name + ".values"][row0:row0 + rows]
            assert np.array_equal(nulls, e_nulls), name
            if ptype == PP.BYTE_ARRAY:
                got = np.array(["" if v is None else v.decode() for v in values])
                assert np.array_equal(got, e_vals), name
            else:
                got = np.asarray(values)
                assert got.dtype.itemsize == e_vals.dtype.itemsize or ptype == PP.BOOLEAN, name
                assert np.array_equal(got.view(e_vals.dtype) if ptype != PP.BOOLEAN else got, e_vals), name  # bit-exact (DOUBLE as bits)
            if info["all_dictionary"]:
                ids = info["ids"]
                assert np.array_equal(ids < 0, e_nulls.astype(bool)), name
            seen.add((name, info["all_dictionary"], tuple(p[0] for p in info["pages"])))
        row0 += rows
    return seen


def test_golden_lineitem_v1():
    seen = check_file(os.path.join(GOLDEN, "lineitem_v1.parquet"), expected_of("lineitem_v1"))
    kinds = {n: d for n, d, _ in seen}
    assert kinds["returnflag"] and kinds["linestatus"] and kinds["discount"] and not kinds["extendedprice"] and not kinds["shipdate"]
    assert any(len(p) > 2 for _, _, p in seen)  # several data pages per chunk


def test_golden_nullable_v2():
    seen = check_file(os.path.join(GOLDEN, "nullable_v2.parquet"), expected_of("nullable_v2"))
    assert any("data_v2" in p for _, _, p in seen)
    assert {n for n, d, _ in seen if d} >= {"runs", "flag"}


@pytest.mark.parametrize("version,page_size,dict_cols", [("1.0", 64, True), ("2.0", 1 << 20, True), ("1.0", 300, False), ("2.0", 97, ["flag", "runs"])])
def test_live_pyarrow_roundtrip(version, page_size, dict_cols):
    pa = pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    sys.path.insert(0, GOLDEN)
    import make_parquet_fixtures as M
    for table in (M.nullable_table(777, seed=21), M.lineitem_table(1501, seed=22)):
        buf = io.BytesIO()
        pq.write_table(table, buf, compression="NONE", data_page_version=version, data_page_size=page_size, use_dictionary=dict_cols,
                       write_statistics=True, row_group_size=500, write_batch_size=37)
        back = pq.read_table(io.BytesIO(buf.getvalue()))      # pyarrow's own reader = the independent implementation
        check_file(buf.getvalue(), M.expected(back))


def test_hybrid_decoder_vectors():
    # parquet-format Encodings.md examples: bit-packed 0..7 at width 3 = 0x88 0xC6 0xFA; RLE run
    buf = bytes([0x03, 0x88, 0xC6, 0xFA])
    assert PP.hybrid_decode(buf, 0, len(buf), 3, 8).tolist() == list(range(8))
    buf = bytes([0x10, 0x05])            # header 16 = RLE run of 8, value 5 in ceil(3/8) = 1 byte
    assert PP.hybrid_decode(buf, 0, len(buf), 3, 8).tolist() == [5] * 8
    buf = bytes([0x06, 0x2C, 0x01, 0x03, 0x88, 0xC6, 0xFA])   # RLE 3 x 300 (width 9 -> 2 bytes), then one bit-packed group... at width 9
    with pytest.raises(PP.Corrupt):
        PP.hybrid_decode(buf, 0, len(buf), 9, 11)            # the packed group needs 9 bytes, only 3 are there
    assert PP.hybrid_decode(buf, 0, 3, 9, 3).tolist() == [300] * 3
    assert PP.hybrid_decode(b"", 0, 0, 0, 0).tolist() == []
    assert PP.hybrid_decode(bytes([0x0A]), 0, 1, 0, 5).tolist() == [0] * 5   # width 0: RLE run without a value byte


def test_rejections():
    rows, cols = chunks_of(os.path.join(GOLDEN, "lineitem_v1.parquet"))[0]
    buf, ptype, max_def, n, tl = cols["quantity"]
    with pytest.raises(PP.Unsupported):
        PP.decode_chunk(buf, ptype, 2, n)
    with pytest.raises(PP.Corrupt):
        PP.decode_chunk(buf[:len(buf) // 2], ptype, max_def, n)
    with pytest.raises(PP.Corrupt):
        PP.decode_chunk(buf, ptype, max_def, n + 5)


def test_dictionary_fallback_to_plain_mid_chunk():
    """a writer whose dictionary page outgrows its limit continues the chunk with PLAIN pages: both kinds in one chunk"""
    pa = pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    rng = np.random.default_rng(4)
    n = 5000
    keys = np.concatenate([rng.integers(0, 20, n // 2), rng.integers(0, 1 << 40, n - n // 2)]).astype(np.int64)
    t = pa.table({"k": pa.array(keys, mask=rng.random(n) < 0.1), "d": pa.array(keys.astype(np.float64) * 0.5)})
    buf = io.BytesIO()
    pq.write_table(t, buf, compression="NONE", use_dictionary=True, dictionary_pagesize_limit=2048, data_page_size=1024, write_batch_size=100,
                   data_page_version="1.0")
    kinds = set()
    for rows, cols in chunks_of(buf.getvalue()):
        for name, (b, pt, md, nv, tl) in cols.items():
            _, _, info = PP.decode_chunk(b, pt, md, nv, tl)
            kinds |= {(p[0], p[2]) for p in info["pages"] if p[0] != "dictionary"}
    assert ("data", PP.ENC_PLAIN) in kinds and (("data", PP.ENC_RLE_DICTIONARY) in kinds or ("data", PP.ENC_PLAIN_DICTIONARY) in kinds)
    sys.path.insert(0, GOLDEN)
    import make_parquet_fixtures as M
    check_file(buf.getvalue(), M.expected(pq.read_table(io.BytesIO(buf.getvalue()))))


def test_hand_built_dictionary_chunk():
    """a chunk assembled byte by byte from the format description (parquet_util.dictionary_chunk): what the oracle must read"""
    ids = [0, 1, 2, 2, 1, 0, 0, 2, 1, 1, 2]
    buf = dictionary_chunk([10, 20, 30], ids, 2)
    values, nulls, info = PP.decode_chunk(buf, PP.INT64, 0, len(ids))
    assert values.tolist() == [[10, 20, 30][i] for i in ids] and not nulls.any() and info["all_dictionary"]
    with pytest.raises(PP.Corrupt):
        PP.decode_chunk(dictionary_chunk([10, 20, 30], [0, 3, 1], 2), PP.INT64, 0, 3)


def decimal_table(n=900, seed=31):
    import decimal
    pa = pytest.importorskip("pyarrow")
    rng = np.random.default_rng(seed)
    D = decimal.Decimal
    short = [None if rng.random() < 0.1 else D(int(rng.integers(-10 ** 11, 10 ** 11))).scaleb(-2) for _ in range(n)]
    few = [D(int(rng.integers(0, 11))).scaleb(-2) for _ in range(n)]
    longs = [None if rng.random() < 0.1 else D(int(rng.integers(-2 ** 62, 2 ** 62)) * int(rng.integers(1, 2 ** 20)) + int(rng.integers(-5, 5))).scaleb(-4) for _ in range(n)]
    small = [D(int(rng.integers(-99999, 99999))).scaleb(-2) for _ in range(n)]
    return pa.table({"price": pa.array(short, pa.decimal128(12, 2)), "discount": pa.array(few, pa.decimal128(12, 2)),
                     "wide": pa.array(longs, pa.decimal128(25, 4)), "small": pa.array(small, pa.decimal128(7, 2))})


def unscaled(col):
    return [None if v is None else int(v.scaleb(-v.as_tuple().exponent)) for v in col.to_pylist()]


@pytest.mark.parametrize("as_integer", [False, True])
def test_decimal_columns(as_integer):
    """DECIMALs as FIXED_LEN_BYTE_ARRAY (big-endian two's complement) and as INT32 / INT64 (parquet-format Logical types): the
    unscaled values must be what pyarrow's reader returns"""
    pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    t = decimal_table()
    buf = io.BytesIO()
    pq.write_table(t, buf, compression="NONE", use_dictionary=["discount"], data_page_size=2048, write_batch_size=64, store_decimal_as_integer=as_integer)
    back = pq.read_table(io.BytesIO(buf.getvalue()))
    kinds = set()
    for rows, cols in chunks_of(buf.getvalue()):
        for name, (b, pt, md, nv, tl) in cols.items():
            values, nulls, info = PP.decode_chunk(b, pt, md, nv, tl)
            want = unscaled(back.column(name))
            assert [None if z else int(v) for v, z in zip(values, nulls)] == want, name
            kinds.add((name, pt))
    if as_integer:
        assert ("small", PP.INT32) in kinds and ("price", PP.INT64) in kinds and ("wide", PP.FIXED_LEN_BYTE_ARRAY) in kinds
    else:
        assert {pt for _, pt in kinds} == {PP.FIXED_LEN_BYTE_ARRAY}
