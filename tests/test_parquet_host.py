"""Host side of the Parquet ingest (csrc/parquet_host.h, through pgpu_parquet_inspect_chunk: no GPU): page headers, dictionary
pages, non-null counts from the definition-level runs and the dictionary / plain classification must agree with the oracle
(oracle/parquet_pages.py, itself pinned against pyarrow) on the golden files and on live pyarrow output; malformed chunks are
rejected, never crash."""
import ctypes
import io
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import parquet_pages as PP  # noqa: E402
from parquet_util import GOLDEN, chunks_of, dictionary_chunk  # noqa: E402
from presto_b200 import _native as N  # noqa: E402


def inspect(buf, ptype, max_def, num_values, tl=0, codec=0):
    c = N.ParquetChunk(bytes(buf), len(buf), ptype, max_def, num_values, tl, codec)
    info = (ctypes.c_int64 * 8)()
    err = ctypes.create_string_buffer(512)
    st = N.lib().pgpu_parquet_inspect_chunk(ctypes.byref(c), info, err, 512)
    return st, list(info), err.value.decode()


RUN = np.dtype([("start", "<i8"), ("src", "<u8"), ("count", "<u4"), ("width", "<u4")])


def runs_of(buf, ptype, max_def, num_values, which, tl=0, codec=0):
    c = N.ParquetChunk(bytes(buf), len(buf), ptype, max_def, num_values, tl, codec)
    n = ctypes.c_int64()
    assert N.lib().pgpu_parquet_chunk_runs(ctypes.byref(c), which, None, 0, ctypes.byref(n)) == N.OK
    out = np.zeros(n.value, RUN)
    if n.value:
        assert N.lib().pgpu_parquet_chunk_runs(ctypes.byref(c), which, out.ctypes.data, n.value, ctypes.byref(n)) == N.OK
    return out


def expand(buf, runs, first, count):
    """numpy statement of the device kernel (parquet.cu k_pq_expand / codeAt): value index -> run by binary search -> code"""
    raw = np.frombuffer(bytes(buf) + b"\0" * 8, np.uint8)
    idx = first + np.arange(count, dtype=np.int64)
    ri = np.searchsorted(runs["start"], idx, side="right") - 1
    assert (ri >= 0).all()
    r = runs[ri]
    assert (idx < r["start"] + r["count"].astype(np.int64)).all(), "run table has gaps"
    width = (r["width"] & 0xFF).astype(np.uint64)
    packed = (r["width"] >> 31) != 0
    bit = r["src"] + (idx - r["start"]).astype(np.uint64) * width
    byte = np.minimum((bit >> np.uint64(3)).astype(np.int64), len(raw) - 8)
    w = np.zeros(count, np.uint64)
    for b in range(5):
        w |= raw[byte + b].astype(np.uint64) << np.uint64(8 * b)
    mask = (np.uint64(1) << width) - np.uint64(1)
    code = (w >> (bit & np.uint64(7))) & mask
    return np.where(packed, code, r["src"]).astype(np.uint32)


def check_runs(name, buf, ptype, max_def, nvals, nulls, o, tl):
    levels = runs_of(buf, ptype, max_def, nvals, 0, tl)
    if nulls.any():
        assert np.array_equal(expand(buf, levels, 0, nvals) == 0, nulls.astype(bool)), name
    else:
        assert len(levels) == 0
    nn = int((nulls == 0).sum())
    values = runs_of(buf, ptype, max_def, nvals, 1, tl)
    pages = runs_of(buf, ptype, max_def, nvals, 2, tl)
    assert int(pages["count"].sum()) == nn
    if o["all_dictionary"]:
        assert np.array_equal(expand(buf, values, 0, nn).astype(np.int32), o["ids"][nulls == 0]), name
    elif ptype == PP.BOOLEAN:
        got = expand(buf, values, 0, nn)
        want, _, _ = PP.decode_chunk(buf, ptype, max_def, nvals, tl)
        assert np.array_equal(got, np.asarray(want)[nulls == 0]), name
    else:  # plain / mixed fixed-width chunk: page by page, as the ingest copies or gathers them
        want, _, _ = PP.decode_chunk(buf, ptype, max_def, nvals, tl)
        dense = np.asarray(want)[nulls == 0]
        dt = dense.dtype
        for pg in pages:
            a, c = int(pg["start"]), int(pg["count"])
            if pg["width"] & 1:
                page_runs = values[(values["start"] >= a) & (values["start"] < a + c)]
                got = np.asarray(o["dictionary"])[expand(buf, page_runs, a, c)]
            else:
                got = np.frombuffer(bytes(buf), dt, c, int(pg["src"]))
            assert np.array_equal(got.view(np.uint8), dense[a:a + c].view(np.uint8)), name


def check_against_oracle(path_or_bytes):
    n = 0
    for rows, cols in chunks_of(path_or_bytes):
        for name, (buf, ptype, max_def, nvals, tl) in cols.items():
            st, info, err = inspect(buf, ptype, max_def, nvals, tl)
            try:
                _, nulls, o = PP.decode_chunk(buf, ptype, max_def, nvals, tl)
            except PP.Unsupported:
                assert st == N.NOT_SUPPORTED, (name, err)
                continue
            if ptype == PP.BYTE_ARRAY and not o["all_dictionary"]:
                assert st == N.NOT_SUPPORTED and "dictionary" in err or "BYTE_ARRAY" in err, (name, err)
                continue
            assert st == N.OK, (name, err)
            data_pages = [p for p in o["pages"] if p[0] != "dictionary"]
            assert info[0] == len(data_pages), name
            assert info[1] == (len(o["dictionary"]) if o["dictionary"] is not None else -1), name
            assert info[2] == int(o["all_dictionary"]), name
            assert info[3] == int((nulls == 0).sum()), name
            assert (info[4] == 0) == (int(nulls.sum()) == 0), name
            assert info[7] == rows
            check_runs(name, buf, ptype, max_def, nvals, nulls, o, tl)
            n += 1
    return n


def test_golden_files():
    assert check_against_oracle(os.path.join(GOLDEN, "lineitem_v1.parquet")) == 16
    assert check_against_oracle(os.path.join(GOLDEN, "nullable_v2.parquet")) == 7


@pytest.mark.parametrize("version,page_size,dict_cols", [("1.0", 64, True), ("2.0", 1 << 20, True), ("1.0", 300, False), ("2.0", 97, ["flag", "runs"])])
def test_live_pyarrow_files(version, page_size, dict_cols):
    pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    sys.path.insert(0, GOLDEN)
    import make_parquet_fixtures as M
    for table in (M.nullable_table(777, seed=21), M.lineitem_table(1501, seed=22)):
        buf = io.BytesIO()
        pq.write_table(table, buf, compression="NONE", data_page_version=version, data_page_size=page_size, use_dictionary=dict_cols,
                       write_statistics=True, row_group_size=500, write_batch_size=37)
        assert check_against_oracle(buf.getvalue()) > 0


def _write(table, compression, version, **kw):
    import pyarrow.parquet as pq
    buf = io.BytesIO()
    pq.write_table(table, buf, compression=compression, data_page_version=version, **kw)
    return buf.getvalue()


@pytest.mark.parametrize("version", ["1.0", "2.0"])
@pytest.mark.parametrize("compression,codec", [("SNAPPY", 1), ("LZ4", 7), ("GZIP", 2), ("ZSTD", 6)])
def test_compressed_pages_are_inflated_on_the_host(compression, codec, version):
    """SNAPPY, GZIP, ZSTD and LZ4_RAW pages (ParquetCompressionUtils.java:43-70): the host inflates the chunk, everything downstream sees the same
    pages as in the uncompressed file — same page / dictionary / run counts, same run tables (up to byte offsets, headers are
    rewritten), and the oracle, which inflates with Arrow's codecs, decodes the same values"""
    pytest.importorskip("pyarrow")
    sys.path.insert(0, GOLDEN)
    import make_parquet_fixtures as M
    n = 0
    for table in (M.nullable_table(3000, seed=7), M.lineitem_table(3000, seed=8)):
        kw = dict(data_page_size=900, use_dictionary=["flag", "runs", "returnflag", "linestatus", "quantity"], write_batch_size=64)
        packed, plain = chunks_of(_write(table, compression, version, **kw)), chunks_of(_write(table, "NONE", version, **kw))
        for (rows, cols), (_, pcols) in zip(packed, plain):
            for name, chunk in cols.items():
                assert len(chunk) == 6 and chunk[5] == codec, (name, chunk[1:])
                buf, ptype, max_def, nvals, tl = chunk[:5]
                pbuf = pcols[name][0]
                st, info, err = inspect(buf, ptype, max_def, nvals, tl, codec)
                pst, pinfo, _ = inspect(pbuf, ptype, max_def, nvals, tl)
                assert (st, info) == (pst, pinfo), (name, err)
                if st != N.OK:
                    continue
                for which in (0, 1, 2):
                    a, b = runs_of(buf, ptype, max_def, nvals, which, tl, codec), runs_of(pbuf, ptype, max_def, nvals, which, tl)
                    assert np.array_equal(a["start"], b["start"]) and np.array_equal(a["count"], b["count"]) and np.array_equal(a["width"], b["width"]), (name, which)
                    rle = (a["width"] >> 31) == 0
                    if which < 2:
                        assert np.array_equal(a["src"][rle], b["src"][rle]), (name, which)      # RLE runs carry their value, not an offset
                got, want = PP.decode_chunk(buf, ptype, max_def, nvals, tl, codec), PP.decode_chunk(pbuf, ptype, max_def, nvals, tl)
                assert np.array_equal(np.asarray(got[1]), np.asarray(want[1])), name
                live = np.asarray(want[1]) == 0
                assert list(np.asarray(got[0], dtype=object)[live]) == list(np.asarray(want[0], dtype=object)[live]), name
                n += 1
    assert n >= 12


def test_hadoop_framed_lz4_pages():
    """the deprecated codec LZ4 (5) = Hadoop's framing around raw LZ4 blocks, several chunks per block, several blocks per page
    (what the reference's decompressFramed reads); no writer here produces it, so the chunk is built by hand"""
    pytest.importorskip("pyarrow")
    from parquet_util import plain_int64_chunk
    rng = np.random.default_rng(4)
    values = np.repeat(rng.integers(-10**12, 10**12, 3000), 5)          # runs: compressible
    packed, plain = plain_int64_chunk(values, True), plain_int64_chunk(values, False)
    assert len(packed) < len(plain)
    st, info, err = inspect(packed, PP.INT64, 0, len(values), 0, 5)
    pst, pinfo, _ = inspect(plain, PP.INT64, 0, len(values))
    assert (st, info) == (pst, pinfo) == (N.OK, pinfo), err
    a, b = runs_of(packed, PP.INT64, 0, len(values), 2, 0, 5), runs_of(plain, PP.INT64, 0, len(values), 2)
    assert np.array_equal(a["start"], b["start"]) and np.array_equal(a["count"], b["count"])
    bad = bytearray(packed)
    bad[40] ^= 0x55
    assert inspect(bytes(bad), PP.INT64, 0, len(values), 0, 5)[0] in (N.INVALID_ARGUMENT, N.OK)
    assert inspect(packed[:len(packed) - 9], PP.INT64, 0, len(values), 0, 5)[0] == N.INVALID_ARGUMENT


def test_other_codecs_and_corrupt_compressed_pages_are_errors():
    pytest.importorskip("pyarrow")
    sys.path.insert(0, GOLDEN)
    import make_parquet_fixtures as M
    table = M.lineitem_table(3000, seed=3)
    for compression in ("BROTLI",):
        rows, cols = chunks_of(_write(table, compression, "1.0"))[0]
        b, pt, md, n, tl, codec = cols["extendedprice"]
        st, _, err = inspect(b, pt, md, n, tl, codec)
        assert st == N.NOT_SUPPORTED and "codec" in err, err
    rng = np.random.default_rng(1)
    for compression in ("SNAPPY", "GZIP", "LZ4", "ZSTD"):
        rows, cols = chunks_of(_write(table, compression, "1.0", use_dictionary=False))[0]
        b, pt, md, n, tl, codec = cols["extendedprice"]
        assert inspect(b, pt, md, n, tl, codec)[0] == N.OK
        assert inspect(b, pt, md, n, tl, 0)[0] == N.NOT_SUPPORTED          # compressed pages handed over as uncompressed
        for trial in range(40):                                            # damaged payloads: an error, never a crash
            bad = bytearray(b)
            for at in rng.integers(40, len(bad), 6):
                bad[at] ^= int(rng.integers(1, 256))
            st, _, err = inspect(bytes(bad), pt, md, n, tl, codec)
            assert st in (N.OK, N.INVALID_ARGUMENT, N.NOT_SUPPORTED), err
            if compression == "GZIP":
                assert st != N.OK, "a damaged gzip member fails its CRC-32"
        st, _, err = inspect(b[:len(b) // 2], pt, md, n, tl, codec)
        assert st == N.INVALID_ARGUMENT


def test_rejections_and_corruption():
    rows, cols = chunks_of(os.path.join(GOLDEN, "nullable_v2.parquet"))[0]
    b, pt, md, n, tl = cols["runs"]
    assert inspect(b, pt, md, n)[0] == N.OK
    st, _, err = inspect(b, pt, 3, n)
    assert st == N.NOT_SUPPORTED and "nested" in err
    st, _, err = inspect(b, 3, md, n)          # INT96
    assert st == N.NOT_SUPPORTED and "physical type" in err
    st, _, err = inspect(b, pt, md, n + 1)
    assert st == N.INVALID_ARGUMENT and "ends after" in err
    st, _, err = inspect(b[:len(b) * 2 // 3], pt, md, n)
    assert st == N.INVALID_ARGUMENT
    st, _, err = inspect(b"", pt, md, 5)
    assert st == N.INVALID_ARGUMENT
    assert inspect(b"", pt, md, 0)[0] == N.OK
    # a data page before the dictionary page's ids can be resolved: drop the dictionary page
    h, body = PP.read_page_header(bytes(b), 0)
    assert h["type"] == PP.PAGE_DICTIONARY
    st, _, err = inspect(b[body + h["compressed_size"]:], pt, md, n)
    assert st == N.INVALID_ARGUMENT and "Dictionary is missing" in err


def test_fuzzed_chunks_never_crash():
    rng = np.random.default_rng(99)
    for fname in ("nullable_v2.parquet", "lineitem_v1.parquet"):
        for rows, cols in chunks_of(os.path.join(GOLDEN, fname)):
            for name, (b, pt, md, n, tl) in cols.items():
                raw = np.frombuffer(b, np.uint8)
                for trial in range(60):
                    x = raw.copy()
                    k = int(rng.integers(1, 4))
                    x[rng.integers(0, min(len(x), 64 if trial % 2 else len(x)), k)] = rng.integers(0, 256, k)
                    cut = len(x) if trial % 3 else int(rng.integers(0, len(x)))
                    st, info, err = inspect(x[:cut].tobytes(), pt, md, n)
                    assert st in (N.OK, N.NOT_SUPPORTED, N.INVALID_ARGUMENT)
                    if st == N.OK:
                        assert 0 <= info[3] <= n and info[7] == n


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
    assert check_against_oracle(buf.getvalue()) > 0


def test_hand_built_dictionary_chunk():
    ids = [0, 1, 2, 2, 1, 0, 0, 2, 1, 1, 2]
    buf = dictionary_chunk([10, 20, 30], ids, 2)
    st, info, err = inspect(buf, PP.INT64, 0, len(ids))
    assert st == N.OK and info[:4] == [1, 3, 1, 11] and info[4] == 0 and info[5] == 1, (info, err)
    assert expand(buf, runs_of(buf, PP.INT64, 0, len(ids), 1), 0, len(ids)).tolist() == ids


def test_plans_of_the_gpu_suite_compile_for_the_signatures_the_ingest_produces():
    """the page signatures pgpu_add_parquet_row_group hands to the kernels (dictionary / plain, with or without a NULL vector)
    for the golden files, compiled for sm_100a without a GPU (also warms the in-tree kernel cache for the GPU box)"""
    from presto_b200 import tpch
    from presto_b200.operators import GpuPipelineOperator
    import test_gpu_parquet as G

    def signature(cols, names):
        enc, nullable = [], []
        for c in names:
            st, info, err = inspect(*cols[c])
            assert st == N.OK, err
            enc.append(1 if info[2] else 0)
            nullable.append(info[4] > 0)
        return enc, nullable

    from presto_b200.operators import plan_text
    from presto_b200.relational import field
    types = [t for _, t in G.NULLABLE]
    jobs = [(plan_text(types, None, [field(i, t) for i, t in enumerate(types)]), "nullable_v2.parquet", [c for c, _ in G.NULLABLE]),
            (tpch.q1_plan(), "lineitem_v1.parquet", ["returnflag", "linestatus", "quantity", "extendedprice", "discount", "tax", "shipdate"]),
            (tpch.q6_plan(), "lineitem_v1.parquet", ["extendedprice", "discount", "shipdate", "quantity"]),
            (tpch.highcard_plan(1000), "lineitem_v1.parquet", ["orderkey", "quantity"])]
    for plan, fname, names in jobs:
        op = GpuPipelineOperator(plan, device=-1)
        try:
            for rows, cols in chunks_of(os.path.join(GOLDEN, fname)):
                enc, nullable = signature(cols, names)
                op.compileSignature(enc, nullable)
            assert op.stats().jit_compiles + op.stats().jit_cache_hits >= 2
        finally:
            op.close()


@pytest.mark.parametrize("as_integer", [False, True])
def test_tpch_decimal_lineitem_file(as_integer):
    """sf0.01 lineitem with DECIMAL(12,2) columns (FIXED_LEN_BYTE_ARRAY(6), or INT64 when stored as integers): the oracle reads the
    generator's rows back, and the product's page / run tables reproduce them"""
    pytest.importorskip("pyarrow")
    from parquet_util import Q1_COLUMNS, tpch_lineitem_parquet
    from presto_b200 import tpch
    data = tpch_lineitem_parquet(0.01, 25_000, as_integer)
    groups = chunks_of(data)
    assert sum(g[0] for g in groups) == 60_175 and len(groups) == 3
    at = 0
    for rows, cols in groups:
        for name in Q1_COLUMNS:
            b, pt, md, nv, tl = cols[name]
            values, nulls, o = PP.decode_chunk(b, pt, md, nv, tl)
            assert not nulls.any()
            if name in ("quantity", "extendedprice", "discount", "tax"):
                assert pt == (PP.INT64 if as_integer else PP.FIXED_LEN_BYTE_ARRAY) and (as_integer or tl == 6)
                want = tpch.tpch_host_column(name, 0.01, 0, tpch.tpch_order_count(0.01), True)[at:at + rows]
                assert np.array_equal(np.asarray(values).astype(np.int64), want), name
                assert o["all_dictionary"] == (name != "extendedprice")
            st, info, err = inspect(b, pt, md, nv, tl)
            assert st == N.OK and info[3] == rows, (name, err)
            if pt != PP.FIXED_LEN_BYTE_ARRAY:
                check_runs(name, b, pt, md, nv, nulls, o, tl)
            else:
                pages = runs_of(b, pt, md, nv, 2, tl)
                assert int(pages["count"].sum()) == rows
                if o["all_dictionary"]:
                    assert np.array_equal(expand(b, runs_of(b, pt, md, nv, 1, tl), 0, rows).astype(np.int32), o["ids"])
                else:  # PLAIN pages: big-endian values at the byte offsets the ingest converts from
                    got = np.concatenate([[int.from_bytes(b[int(pg["src"]) + i * tl:int(pg["src"]) + (i + 1) * tl], "big", signed=True) for i in range(int(pg["count"]))]
                                          for pg in pages])
                    assert np.array_equal(got.astype(np.int64), np.asarray(values).astype(np.int64))
        at += rows


def test_fixed_len_byte_array_limits():
    b = b"\x00" * 40
    st, _, err = inspect(b, PP.FIXED_LEN_BYTE_ARRAY, 0, 1, 0)
    assert st == N.NOT_SUPPORTED and "FIXED_LEN_BYTE_ARRAY" in err
    st, _, err = inspect(b, PP.FIXED_LEN_BYTE_ARRAY, 0, 1, 17)
    assert st == N.NOT_SUPPORTED


def test_page_parallel_walk_gives_the_same_tables(monkeypatch):
    """a chunk of 200 small pages walked by 1 and by 6 host threads (PGPU_PARQUET_THREADS): identical run and page tables"""
    pytest.importorskip("pyarrow")
    import pyarrow as pa
    import pyarrow.parquet as pq
    rng = np.random.default_rng(12)
    n = 40_000
    t = pa.table({"k": pa.array(rng.integers(0, 9, n).astype(np.int32), mask=rng.random(n) < 0.2), "b": pa.array(rng.random(n) < 0.3, mask=rng.random(n) < 0.1)})
    buf = io.BytesIO()
    pq.write_table(t, buf, compression="NONE", use_dictionary=["k"], data_page_size=128, write_batch_size=200, row_group_size=n)
    (rows, cols), = chunks_of(buf.getvalue())
    for name, (b, pt, md, nv, tl) in cols.items():
        tables = []
        for threads in ("1", "6"):
            monkeypatch.setenv("PGPU_PARQUET_THREADS", threads)
            st, info, err = inspect(b, pt, md, nv, tl)
            assert st == N.OK and info[0] >= (64 if name == "k" else 30), (name, info, err)   # >= 64 pages: really walked by 2+ threads
            tables.append([runs_of(b, pt, md, nv, which, tl) for which in (0, 1, 2)])
        for a, c in zip(*tables):
            assert np.array_equal(a, c), name
        _, nulls, o = PP.decode_chunk(b, pt, md, nv, tl)
        check_runs(name, b, pt, md, nv, nulls, o, tl)


def test_host_parser_under_sanitizers(tmp_path):
    """csrc/parquet_host.h compiled with AddressSanitizer + UBSan (tests/emu/parquet_host_fuzz.cpp): 150 mutations / truncations of
    every golden chunk, parsed with 1 and 4 page threads from an exactly sized heap copy — no out-of-bounds access, and every
    accepted chunk's run / page tables stay inside the chunk (what the device kernels will read)"""
    import shutil
    import subprocess
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    here = os.path.dirname(os.path.abspath(__file__))
    exe = str(tmp_path / "parquet_host_fuzz")
    r = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-sanitize-recover=all", "-pthread", "-o", exe,
                        os.path.join(here, "emu", "parquet_host_fuzz.cpp")], capture_output=True, text=True)
    if r.returncode != 0 and "asan" in r.stderr.lower():
        pytest.skip("sanitizer runtime not installed")
    assert r.returncode == 0, r.stderr
    accepted = rejected = 0
    for fname in ("nullable_v2.parquet", "lineitem_v1.parquet"):
        for rows, cols in chunks_of(os.path.join(GOLDEN, fname)):
            for name, (b, pt, md, nv, tl) in cols.items():
                path = tmp_path / "chunk.bin"
                path.write_bytes(b)
                r = subprocess.run([exe, str(path), str(pt), str(md), str(nv), str(tl), "150", "11"], capture_output=True, text=True)
                assert r.returncode == 0, (fname, name, r.returncode, r.stderr[-2000:])
                a, j = map(int, r.stdout.split())
                accepted += a
                rejected += j
    assert accepted > 1000 and rejected > 1000
