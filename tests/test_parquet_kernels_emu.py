"""The CUDA kernels of the Parquet ingest (presto_b200/csrc/parquet.cu) executed on the CPU, thread by thread, from the same source
(tests/emu/parquet_kernels_emu.cpp compiles the .cu as C++ with blockIdx / threadIdx supplied), driven in the order
pgpu_add_parquet_row_group drives them, over the product's own run / page tables — against the oracle (oracle/parquet_pages.py).
This checks the kernels' logic without a GPU; tests/test_gpu_parquet.py checks the real execution."""
import ctypes
import io
import os
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
import parquet_pages as PP  # noqa: E402
from parquet_util import GOLDEN, chunks_of  # noqa: E402
from test_parquet_host import inspect, runs_of  # noqa: E402

_emu = None


def emu():
    global _emu
    if _emu is None:
        src = os.path.join(HERE, "emu", "parquet_kernels_emu.cpp")
        cu = os.path.join(os.path.dirname(HERE), "presto_b200", "csrc", "parquet.cu")
        out = os.path.join(HERE, "emu", "_build", "libpqemu.so")
        os.makedirs(os.path.dirname(out), exist_ok=True)
        if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(cu), os.path.getmtime(cu[:-3] + ".h")):
            cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
            subprocess.run(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-w", "-x", "c++", "-I" + os.path.join(cuda, "include"), "-o", out, src], check=True)
        _emu = ctypes.CDLL(out)
    return _emu


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def expand_runs(chunk, runs, first, count, out_width, dictionary, limit):
    out = np.zeros(max(count, 1) * out_width, np.uint8)
    err = np.zeros(1, np.uint32)
    rc = emu().emu_expand_runs(_ptr(chunk), ctypes.c_size_t(len(chunk) - 16), _ptr(runs), ctypes.c_int64(len(runs)), ctypes.c_int64(first), ctypes.c_int64(count),
                               out_width, _ptr(dictionary) if dictionary is not None else None, ctypes.c_uint32(limit), _ptr(out), _ptr(err))
    assert rc == 0
    return out[:count * out_width], int(err[0])


def convert_plain(chunk, off, in_width, kind, count, out_width):
    out = np.zeros(max(count, 1) * out_width, np.uint8)
    src = chunk[off:]
    assert emu().emu_convert_plain(_ptr(src), in_width, kind, ctypes.c_int64(count), _ptr(out), out_width) == 0
    return out[:count * out_width]


def block_ints(raw, width):
    """block bytes -> Python ints (int64, or the sign-magnitude 128-bit words of Int128ArrayBlock)"""
    if width == 8:
        return [int(v) for v in raw.view("<i8")]
    w = raw.view("<u8").reshape(-1, 2)
    return [(-1 if int(hi) >> 63 else 1) * (int(lo) | ((int(hi) & (2 ** 63 - 1)) << 64)) for lo, hi in w]


def decode_with_kernels(buf, ptype, max_def, nvals, tl, out_width):
    """the decode sequence of pgpu_add_parquet_row_group (csrc/pipeline.cpp) for one chunk -> (kind, payload, nulls)"""
    st, info, err = inspect(buf, ptype, max_def, nvals, tl)
    assert st == 0, err
    chunk = np.frombuffer(bytes(buf) + b"\0" * 16, np.uint8).copy()
    levels, vruns, pages = (np.ascontiguousarray(runs_of(buf, ptype, max_def, nvals, w, tl)) for w in (0, 1, 2))
    non_null, dict_count, all_dict = info[3], info[1], bool(info[2])
    nulls = np.zeros(nvals, np.uint8)
    if len(levels):
        bits = np.zeros((nvals + 7) // 8 + 1, np.uint8)
        assert emu().emu_levels_to_null_bits(_ptr(chunk), ctypes.c_size_t(len(buf)), _ptr(levels), ctypes.c_int64(len(levels)), ctypes.c_int64(nvals), _ptr(bits)) == 0
        nulls = np.unpackbits(bits, bitorder="big")[:nvals]          # the wire format's bit order: most significant bit first
    in_width = {PP.INT32: 4, PP.INT64: 8, PP.DOUBLE: 8, PP.BOOLEAN: 1, PP.FIXED_LEN_BYTE_ARRAY: tl}.get(ptype, 0)
    conv = 1 if ptype == PP.FIXED_LEN_BYTE_ARRAY else 0 if (ptype == PP.INT32 and out_width == 8) else -1
    dict_off = 0
    if dict_count >= 0:
        h, dict_off = PP.read_page_header(bytes(buf), 0)
        assert h["type"] == PP.PAGE_DICTIONARY

    def plain(off, count):
        if conv < 0:
            return chunk[off:off + count * out_width].copy()
        return convert_plain(chunk, off, in_width, conv, count, out_width)
    if all_dict:
        ids, e = expand_runs(chunk, vruns, 0, non_null, 4, None, dict_count)
        assert e == 0
        full = np.zeros(nvals, np.int32)
        full[nulls == 0] = ids.view("<i4")                              # expandNulls (serde.cu, proven on the GPU)
        if nulls.any():
            assert emu().emu_null_ids(_ptr(full), _ptr(nulls), ctypes.c_int64(nvals), dict_count) == 0
        dictionary = plain(dict_off, dict_count) if in_width else None
        return "dictionary", (full, dictionary), nulls
    if ptype == PP.BOOLEAN:
        dense, e = expand_runs(chunk, vruns, 0, non_null, 1, None, 2)
        assert e == 0
        return "plain", dense, nulls
    dictionary = plain(dict_off, dict_count) if info[6] else None
    dense = np.zeros(non_null * out_width, np.uint8)
    for pg in pages:
        a, c = int(pg["start"]), int(pg["count"])
        if not c:
            continue
        if pg["width"] & 1:
            page_runs = np.ascontiguousarray(vruns[(vruns["start"] >= a) & (vruns["start"] < a + c)])
            got, e = expand_runs(chunk, page_runs, a, c, out_width, dictionary, dict_count)
            assert e == 0
        else:
            got = plain(int(pg["src"]), c)
        dense[a * out_width:(a + c) * out_width] = got
    return "plain", dense, nulls


def check_chunk(name, buf, ptype, max_def, nvals, tl, widen=False):
    if ptype == PP.BYTE_ARRAY:
        out_width = 0
    elif ptype == PP.FIXED_LEN_BYTE_ARRAY:
        out_width = 8 if tl <= 8 else 16
    else:
        out_width = {PP.INT32: 8 if widen else 4, PP.INT64: 8, PP.DOUBLE: 8, PP.BOOLEAN: 1}[ptype]
    want, want_nulls, o = PP.decode_chunk(buf, ptype, max_def, nvals, tl)
    kind, payload, nulls = decode_with_kernels(buf, ptype, max_def, nvals, tl, out_width)
    assert np.array_equal(nulls, want_nulls), name
    present = want_nulls == 0
    as_int = ptype == PP.FIXED_LEN_BYTE_ARRAY or (ptype == PP.INT32 and widen)
    if kind == "dictionary":
        ids, dictionary = payload
        assert o["all_dictionary"]
        assert np.array_equal(ids[present], o["ids"][present]) and (ids[~present] == len(o["dictionary"])).all(), name
        if ptype != PP.BYTE_ARRAY:
            got = block_ints(dictionary, out_width) if as_int else dictionary
            exp = [int(v) for v in o["dictionary"]] if as_int else np.asarray(o["dictionary"]).view(np.uint8)
            assert list(got) == list(exp), name
        return kind
    dense = payload
    if ptype == PP.BOOLEAN:
        assert np.array_equal(dense, np.asarray(want)[present]), name
    elif as_int:
        assert block_ints(dense, out_width) == [int(v) for v in np.asarray(want)[present]], name
    else:
        assert np.array_equal(dense, np.ascontiguousarray(np.asarray(want)[present]).view(np.uint8)), name
    return kind


def check_file(data, widen=False):
    kinds = set()
    for rows, cols in chunks_of(data):
        for name, (b, pt, md, nv, tl) in cols.items():
            try:
                PP.decode_chunk(b, pt, md, nv, tl)
            except PP.Unsupported:
                continue
            if pt == PP.BYTE_ARRAY and inspect(b, pt, md, nv, tl)[0] != 0:
                continue
            kinds.add((name, check_chunk(name, b, pt, md, nv, tl, widen)))
    return kinds


def test_golden_files_through_the_kernels():
    kinds = check_file(os.path.join(GOLDEN, "nullable_v2.parquet"))
    assert {("runs", "dictionary"), ("flag", "dictionary"), ("allnull", "dictionary"), ("x", "plain"), ("b", "plain"), ("k", "plain"), ("big", "plain")} <= kinds
    kinds = check_file(os.path.join(GOLDEN, "lineitem_v1.parquet"))
    assert ("returnflag", "dictionary") in kinds and ("extendedprice", "plain") in kinds and ("shipdate", "plain") in kinds
    check_file(os.path.join(GOLDEN, "lineitem_v1.parquet"), widen=True)      # INT32 ship dates widened to 64 bits


@pytest.mark.parametrize("version,page_size,dict_cols", [("1.0", 64, True), ("2.0", 1 << 20, True), ("1.0", 300, False), ("2.0", 97, ["flag", "runs"])])
def test_live_pyarrow_files_through_the_kernels(version, page_size, dict_cols):
    pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    sys.path.insert(0, GOLDEN)
    import make_parquet_fixtures as M
    for table in (M.nullable_table(777, seed=21), M.lineitem_table(1501, seed=22)):
        buf = io.BytesIO()
        pq.write_table(table, buf, compression="NONE", data_page_version=version, data_page_size=page_size, use_dictionary=dict_cols, row_group_size=500,
                       write_batch_size=37)
        assert check_file(buf.getvalue())


@pytest.mark.parametrize("as_integer", [False, True])
def test_decimal_files_through_the_kernels(as_integer):
    """big-endian FIXED_LEN_BYTE_ARRAY decimals -> int64 / sign-magnitude int128 (k_pq_convert), INT32 decimals widened"""
    pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    import test_parquet_oracle as O
    buf = io.BytesIO()
    pq.write_table(O.decimal_table(), buf, compression="NONE", use_dictionary=["discount"], data_page_size=2048, write_batch_size=64, store_decimal_as_integer=as_integer)
    kinds = check_file(buf.getvalue(), widen=True)
    assert ("discount", "dictionary") in kinds and ("wide", "plain") in kinds


def test_dictionary_fallback_and_bad_ids_through_the_kernels():
    pa = pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    from parquet_util import dictionary_chunk
    rng = np.random.default_rng(4)
    n = 5000
    keys = np.concatenate([rng.integers(0, 20, n // 2), rng.integers(0, 1 << 40, n - n // 2)]).astype(np.int64)
    t = pa.table({"k": pa.array(keys, mask=rng.random(n) < 0.1)})
    buf = io.BytesIO()
    pq.write_table(t, buf, compression="NONE", use_dictionary=True, dictionary_pagesize_limit=2048, data_page_size=1024, write_batch_size=100, data_page_version="1.0")
    assert check_file(buf.getvalue()) == {("k", "plain")}          # mixed chunk: dictionary pages gathered, PLAIN pages copied
    # an id past the dictionary raises the device error flag (kPqErrBadId = 64) and is stored as 0
    bad = dictionary_chunk([10, 20, 30], [0, 1, 2, 3, 1, 0, 0, 2, 1, 1, 2], 2)
    chunk = np.frombuffer(bad + b"\0" * 16, np.uint8).copy()
    runs = np.ascontiguousarray(runs_of(bad, PP.INT64, 0, 11, 1))
    ids, err = expand_runs(chunk, runs, 0, 11, 4, None, 3)
    assert err == 64 and ids.view("<i4").tolist() == [0, 1, 2, 0, 1, 0, 0, 2, 1, 1, 2]
