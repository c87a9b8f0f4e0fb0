"""The host-side decoders of compressed Parquet pages and LZ4 SerializedPages (presto_b200/csrc/parquet_host.h, zstd_decode*.h:
Snappy, LZ4 block / Hadoop frames, gzip, Zstandard) against Arrow's encoders, byte for byte, on data of every texture: runs,
text, random bytes, numeric columns, empty and one-byte inputs, inputs of several zstd blocks. tests/emu/codec_test.cpp is built
with AddressSanitizer + UBSan: truncated and bit-flipped inputs must be refused (or decode to other bytes), never fault."""
import os
import subprocess

import numpy as np
import pytest

pa = pytest.importorskip("pyarrow")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def binary(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("codec") / "codec_test")
    subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-pthread", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-I", os.path.join(ROOT, "presto_b200", "csrc"),
                    os.path.join(ROOT, "tests", "emu", "codec_test.cpp"), "-o", out], check=True)
    return out


def samples():
    rng = np.random.default_rng(12)
    words = [b"lineitem", b"orderkey", b" ", b"0.05", b"AIR", b"DELIVER IN PERSON", b"\n", b"1996-03-13", b"N", b"O", b"fluffily regular"]
    out = {
        "empty": b"",
        "one_byte": b"x",
        "zeros": bytes(200_000),
        "runs": np.repeat(rng.integers(0, 256, 3000).astype(np.uint8), rng.integers(1, 200, 3000)).tobytes(),
        "text": b"".join(words[int(i)] for i in rng.integers(0, len(words), 60_000)),
        "random": rng.integers(0, 256, 300_000).astype(np.uint8).tobytes(),
        "doubles": (rng.integers(100, 10_000_000, 80_000) / 100.0).astype("<f8").tobytes(),
        "int64_sorted": np.cumsum(rng.integers(0, 5, 120_000)).astype("<i8").tobytes(),
        "dictionary_ids": rng.integers(0, 7, 500_000).astype(np.uint8).tobytes(),
        "mixed_large": b"".join([rng.integers(0, 4, 150_000).astype(np.uint8).tobytes(), rng.integers(0, 256, 150_000).astype(np.uint8).tobytes(),
                                 b"abcdefgh" * 40_000, np.arange(60_000, dtype="<i4").tobytes()]),
    }
    return out


def encode(codec, raw, level=None):
    if codec == "lz4_hadoop":
        import sys
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from parquet_util import lz4_hadoop_frame
        return lz4_hadoop_frame(raw, blocks=3, chunks_per_block=2) if len(raw) >= 6 else lz4_hadoop_frame(raw, 1, 1)
    c = pa.Codec(codec, compression_level=level) if level is not None else pa.Codec(codec)
    return c.compress(raw, asbytes=True)


def run(binary, tmp_path, codec, packed, size):
    src, dst = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    open(src, "wb").write(packed)
    if os.path.exists(dst):
        os.remove(dst)
    r = subprocess.run([binary, codec, src, str(size), dst], stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=120)
    assert r.returncode in (0, 3), r.stderr.decode()[-2000:]      # anything else is a sanitizer report or a crash
    return open(dst, "rb").read() if r.returncode == 0 else None


CODECS = [("snappy", None), ("lz4_raw", None), ("lz4_hadoop", None), ("gzip", 1), ("gzip", 9), ("zstd", 1), ("zstd", 3), ("zstd", 9), ("zstd", 19)]


@pytest.mark.parametrize("codec,level", CODECS, ids=["%s%s" % (c, "" if l is None else "_l%d" % l) for c, l in CODECS])
def test_decoders_equal_arrow_byte_for_byte(binary, tmp_path, codec, level):
    for name, raw in samples().items():
        if codec == "lz4_hadoop" and not raw:
            continue
        packed = encode(codec, raw, level)
        assert run(binary, tmp_path, codec, packed, len(raw)) == raw, (codec, level, name)
        if raw:
            assert run(binary, tmp_path, codec, packed, len(raw) + 1) is None, (name, "a larger declared size must be refused")
            assert run(binary, tmp_path, codec, packed, len(raw) - 1) is None, (name, "a smaller declared size must be refused")


@pytest.mark.parametrize("codec", ["snappy", "lz4_raw", "lz4_hadoop", "gzip", "zstd"])
def test_damaged_inputs_never_fault(binary, tmp_path, codec):
    rng = np.random.default_rng(5)
    s = samples()
    for name in ("text", "runs", "doubles", "mixed_large"):
        raw = s[name]
        packed = encode(codec, raw, 3 if codec == "zstd" else None)
        for trial in range(12):
            bad = bytearray(packed)
            if trial % 3 == 0:
                bad = bad[:int(rng.integers(0, len(bad)))]
            else:
                for at in rng.integers(0, len(bad), int(rng.integers(1, 6))):
                    bad[at] ^= 1 << int(rng.integers(0, 8))
            got = run(binary, tmp_path, codec, bytes(bad), len(raw))      # asserts "no sanitizer report" inside
            if codec == "gzip" and bytes(bad) != packed:
                assert got is None or got == raw                          # CRC-32: damage is detected (or did not touch the stream)
