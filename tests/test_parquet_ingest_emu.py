"""The device side of pgpu_add_parquet_row_group on the CPU: the product's orchestration (csrc/parquet_ingest.cpp — arena layout,
copies, launch order) and kernels (csrc/parquet.cu) compiled as C++ with memory copies and kernel launches emulated
(tests/emu/parquet_ingest_emu.cpp). The Blocks it produces — plain values + null vectors, or ids + dictionary with the NULL entry —
must be what the oracle / pyarrow read from the same chunks."""
import ctypes
import io
import os
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import parquet_pages as PP  # noqa: E402
from parquet_util import GOLDEN, Q1_COLUMNS, chunks_of  # noqa: E402
from presto_b200 import _native as N  # noqa: E402
from presto_b200 import tpch  # noqa: E402
from presto_b200.operators import plan_text  # noqa: E402
from presto_b200.relational import field  # noqa: E402
from presto_b200.spi import BIGINT, BOOLEAN, DATE, DOUBLE, INTEGER, createDecimalType, createVarcharType  # noqa: E402

_lib = None


def lib():
    global _lib
    if _lib is None:
        csrc = os.path.join(ROOT, "presto_b200", "csrc")
        src = os.path.join(HERE, "emu", "parquet_ingest_emu.cpp")
        out = os.path.join(HERE, "emu", "_build", "libpqingest.so")
        deps = [src] + [os.path.join(csrc, f) for f in ("parquet_ingest.cpp", "parquet_ingest.h", "parquet.cu", "parquet.h", "parquet_host.h", "serde.h", "plan.cpp", "plan.h")]
        os.makedirs(os.path.dirname(out), exist_ok=True)
        if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
            cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
            subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-shared", "-fPIC", "-w", "-pthread", "-x", "c++", "-I" + os.path.join(cuda, "include"), "-o", out, src,
                            os.path.join(csrc, "plan.cpp")], check=True)
        _lib = ctypes.CDLL(out)
        _lib.emu_decode_row_group.restype = ctypes.c_void_p
        _lib.emu_decode_row_group.argtypes = [ctypes.c_char_p, ctypes.POINTER(N.ParquetChunk), ctypes.c_int32, ctypes.c_int64, ctypes.c_int32, ctypes.POINTER(ctypes.c_int32),
                                              ctypes.c_char_p, ctypes.c_size_t]
        _lib.emu_column.restype = ctypes.POINTER(N.Column)
        _lib.emu_column.argtypes = [ctypes.c_void_p, ctypes.c_int32]
        _lib.emu_error_bits.argtypes = [ctypes.c_void_p]
        _lib.emu_error_bits.restype = ctypes.c_uint32
        _lib.emu_free.argtypes = [ctypes.c_void_p]
    return _lib


def _array(ptr, dtype, count):
    if not ptr or count == 0:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ctypes.cast(ptr, ctypes.POINTER(ctypes.c_uint8)), shape=(count * np.dtype(dtype).itemsize,)).view(dtype).copy()


def _ints(raw_u8, width):
    if width == 16:
        w = raw_u8.view("<u8").reshape(-1, 2)
        return [(-1 if int(hi) >> 63 else 1) * (int(lo) | ((int(hi) & (2 ** 63 - 1)) << 64)) for lo, hi in w]
    return [int(v) for v in raw_u8.view({1: "i1", 4: "<i4", 8: "<i8"}[width])]


def decode(types, chunks, rows, threads=1):
    """-> per channel (values as Python list with None for NULL, was it a DictionaryBlock), error bits"""
    plan = plan_text(types, None, [field(i, t) for i, t in enumerate(types)]).encode()
    arr = (N.ParquetChunk * len(chunks))()
    keep = []
    for i, ch in enumerate(chunks):
        data = bytes(ch[0])
        keep.append(data)
        arr[i].data, arr[i].len, arr[i].physical_type, arr[i].max_definition_level, arr[i].num_values, arr[i].type_length = data, len(data), ch[1], ch[2], ch[3], ch[4]
    status = ctypes.c_int32()
    msg = ctypes.create_string_buffer(512)
    s = lib().emu_decode_row_group(plan, arr, len(chunks), rows, threads, ctypes.byref(status), msg, 512)
    if not s:
        return status.value, msg.value.decode()
    out = []
    for c, t in enumerate(types):
        col = lib().emu_column(s, c).contents
        assert col.position_count == rows and col.type_kind == t.kind
        width = 0 if t.kind in (6, 7) else (1 if t.kind == 0 else 4 if t.kind in (2, 3) else 16 if (t.kind == 5 and t.precision > 18) else 8)

        def values_of(column, count):
            if width == 0:
                off = _array(column.offsets, "<i4", count + 1)
                data = _array(column.values, np.uint8, int(off[-1]) if count else 0).tobytes()
                return [data[off[k]:off[k + 1]].decode() for k in range(count)]
            raw = _array(column.values, np.uint8, count * width)
            if t.kind == 4:
                return [float(v) for v in raw.view("<f8")]
            return _ints(raw, width)
        if col.encoding == N.ENC_DICTIONARY:
            d = col.dictionary.contents
            entries = values_of(d, d.position_count)
            dnulls = _array(d.nulls, np.uint8, d.position_count) if d.nulls else np.zeros(d.position_count, np.uint8)
            ids = _array(col.ids, "<i4", rows)
            assert ids.min(initial=0) >= 0 and ids.max(initial=0) < d.position_count
            out.append(([None if dnulls[k] else entries[k] for k in ids], True))
        else:
            assert col.encoding == N.ENC_PLAIN
            vals = values_of(col, rows)
            nulls = _array(col.nulls, np.uint8, rows) if col.nulls else np.zeros(rows, np.uint8)
            out.append(([None if z else v for v, z in zip(vals, nulls)], False))
    err = lib().emu_error_bits(s)
    lib().emu_free(s)
    return out, err


def expected_column(chunk, t):
    b, pt, md, nv, tl = chunk
    values, nulls, info = PP.decode_chunk(b, pt, md, nv, tl)
    if pt == PP.BYTE_ARRAY:
        n = t.precision
        return [None if z else (v.decode()[:n] if n else v.decode()) for v, z in zip(values, nulls)]
    if t.kind == 4:
        return [None if z else float(v) for v, z in zip(values, nulls)]
    return [None if z else int(v) for v, z in zip(values, nulls)]


def check(types, names, data, threads=1):
    seen = set()
    for rows, cols in chunks_of(data):
        chunks = [cols[n] for n in names]
        got, err = decode(types, chunks, rows, threads)
        assert err == 0
        for (vals, is_dict), name, t, ch in zip(got, names, types, chunks):
            want = expected_column(ch, t)
            if t.kind == 4:   # DOUBLE: compare bit patterns
                same = [a is None and b is None or (a is not None and b is not None and np.float64(a).view(np.int64) == np.float64(b).view(np.int64)) for a, b in zip(vals, want)]
                assert all(same), name
            else:
                assert vals == want, name
            seen.add((name, is_dict))
    return seen


V1, V3 = createVarcharType(1), createVarcharType(3)


def test_golden_nullable_file():
    types = [BIGINT, INTEGER, DOUBLE, V3, BOOLEAN, BIGINT, BIGINT]
    names = ["k", "runs", "x", "flag", "b", "allnull", "big"]
    seen = check(types, names, os.path.join(GOLDEN, "nullable_v2.parquet"))
    assert {("runs", True), ("flag", True), ("allnull", True), ("k", False), ("x", False), ("b", False), ("big", False)} == seen


def test_golden_lineitem_file_and_varchar_truncation():
    names = Q1_COLUMNS + ["orderkey"]
    types = [V1, V1, DOUBLE, DOUBLE, DOUBLE, DOUBLE, DATE, BIGINT]
    seen = check(types, names, os.path.join(GOLDEN, "lineitem_v1.parquet"))
    assert ("returnflag", True) in seen and ("extendedprice", False) in seen
    # INT32 ship dates into a BIGINT channel (widened), as IntColumnReader allows
    check([BIGINT], ["shipdate"], os.path.join(GOLDEN, "lineitem_v1.parquet"))
    # a VARCHAR(1) channel truncates 'bb' / 'ccc' to one code point (BinaryColumnReader.java:44-58)
    check([V1], ["flag"], os.path.join(GOLDEN, "nullable_v2.parquet"))


@pytest.mark.parametrize("as_integer", [False, True])
def test_decimal_columns(as_integer):
    pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    import test_parquet_oracle as O
    buf = io.BytesIO()
    pq.write_table(O.decimal_table(), buf, compression="NONE", use_dictionary=["discount"], data_page_size=2048, write_batch_size=64, store_decimal_as_integer=as_integer)
    types = [createDecimalType(12, 2), createDecimalType(12, 2), createDecimalType(25, 4), createDecimalType(7, 2)]
    seen = check(types, ["price", "discount", "wide", "small"], buf.getvalue())
    assert ("discount", True) in seen and ("wide", False) in seen


def test_tpch_decimal_lineitem_and_page_threads():
    pytest.importorskip("pyarrow")
    from parquet_util import tpch_lineitem_parquet
    d = createDecimalType(12, 2)
    types = [V1, V1, d, d, d, d, DATE]
    data = tpch_lineitem_parquet(0.01, 25_000)
    assert check(types, Q1_COLUMNS, data) == check(types, Q1_COLUMNS, data, threads=4)


def test_dictionary_fallback_and_live_files():
    pa = pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    rng = np.random.default_rng(4)
    n = 5000
    keys = np.concatenate([rng.integers(0, 20, n // 2), rng.integers(0, 1 << 40, n - n // 2)]).astype(np.int64)
    t = pa.table({"k": pa.array(keys, mask=rng.random(n) < 0.1), "d": pa.array(keys.astype(np.float64) * 0.5)})
    buf = io.BytesIO()
    pq.write_table(t, buf, compression="NONE", use_dictionary=True, dictionary_pagesize_limit=2048, data_page_size=1024, write_batch_size=100, data_page_version="1.0")
    assert check([BIGINT, DOUBLE], ["k", "d"], buf.getvalue()) == {("k", False), ("d", False)}
    sys.path.insert(0, GOLDEN)
    import make_parquet_fixtures as M
    for version, page_size in (("1.0", 64), ("2.0", 97)):
        buf = io.BytesIO()
        pq.write_table(M.nullable_table(777, seed=21), buf, compression="NONE", data_page_version=version, data_page_size=page_size, use_dictionary=["flag", "runs", "allnull"],
                       row_group_size=500, write_batch_size=37)
        check([BIGINT, INTEGER, DOUBLE, V3, BOOLEAN, BIGINT, BIGINT], ["k", "runs", "x", "flag", "b", "allnull", "big"], buf.getvalue())


def test_rejections_and_bad_ids():
    from parquet_util import dictionary_chunk
    rows, cols = chunks_of(os.path.join(GOLDEN, "lineitem_v1.parquet"))[0]
    st, msg = decode([DOUBLE], [cols["orderkey"]], rows)
    assert st == N.NOT_SUPPORTED and "physical type" in msg
    st, msg = decode([DOUBLE, DOUBLE], [cols["quantity"]], rows)
    assert st == N.INVALID_ARGUMENT and "column chunks" in msg
    b, pt, md, nv, tl = cols["extendedprice"]
    st, msg = decode([DOUBLE], [(b[:len(b) // 2], pt, md, nv, tl)], rows)
    assert st == N.INVALID_ARGUMENT and "channel 0" in msg
    bad = dictionary_chunk([10, 20, 30], [0, 1, 2, 3, 1, 0, 0, 2, 1, 1, 2], 2)
    got, err = decode([BIGINT], [(bad, 2, 0, 11, 0)], 11)
    assert err == 64 and got[0][0] == [10, 20, 30, 10, 20, 10, 10, 30, 20, 20, 30]
