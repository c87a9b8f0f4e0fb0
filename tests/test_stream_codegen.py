"""Streaming runs, the host-checkable part (no GPU): which plans get the kernels, and the generated `P::key_order` — the order the
run buffer's rows are expected in — compiled for the host and checked against Python's ordering of the same tuples. The kernels
themselves are covered by tests/test_gpu_stream.py."""
import ctypes
import os
import struct
import subprocess

import numpy as np
import pytest

from presto_b200.operators import Aggregation, GpuPipelineOperator, plan_text
from presto_b200.relational import compare, constant, field
from presto_b200.spi import BIGINT, DOUBLE, INTEGER

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "emu", "_build")
NO_DIRECTORY = (1 << 20) + 1

KERNELS = ("pgpu_k_runs_count", "pgpu_k_runs_write", "pgpu_k_runs_emit", "pgpu_k_runs_to_table")


def source_of(plan, ncols):
    op = GpuPipelineOperator(plan, -1)          # device -1: compile only
    try:
        op.compileSignature([0] * ncols, [False] * ncols)
        return op.generatedSource()
    finally:
        op.close()


def group_plan(step="single", expected=NO_DIRECTORY, max_partial=0, flt=None):
    proj = None if flt is None else [field(0, BIGINT), field(1, DOUBLE)]
    return plan_text([BIGINT, DOUBLE], flt, proj, step=step, groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)],
                     expectedGroups=expected, maxPartialGroups=max_partial)


def test_streaming_kernels_only_where_the_path_is_eligible(monkeypatch):
    src = source_of(group_plan(), 2)
    assert all(("void __launch_bounds__" in src) and (k + "(") in src for k in KERNELS)
    assert "static int key_order(" in src
    src = source_of(group_plan(flt=compare("lt", field(1, DOUBLE), constant(30.0, DOUBLE))), 2)
    assert all((k + "(") in src for k in KERNELS)                        # a filter is no obstacle: rejected rows form dead runs
    for plan in (group_plan(step="partial"),                             # intermediate output goes through the exchange's table kernels
                 group_plan(step="partial", max_partial=100_000),       # partial flushes
                 group_plan(expected=1000)):                             # few groups: the shared-memory directory path
        src = source_of(plan, 2)
        assert not any((k + "(") in src for k in KERNELS), plan
    monkeypatch.setenv("PGPU_TUNE_STREAM", "0")
    src = source_of(group_plan(), 2)
    assert not any((k + "(") in src for k in KERNELS)


# ------------------------------------------------------------------------------------------------- key_order on the host
def key_order_lib(src, tag):
    """the generated function's text, compiled as it is behind a C entry point"""
    a = src.index("PGPU_DEV static int key_order(")
    b = src.index("\n    }\n", a) + len("\n    }\n")
    os.makedirs(BUILD, exist_ok=True)
    cpp, so = os.path.join(BUILD, "key_order_%s.cpp" % tag), os.path.join(BUILD, "key_order_%s.so" % tag)
    with open(cpp, "w") as f:
        f.write("#include <cstdint>\ntypedef uint64_t u64; typedef int64_t i64; typedef uint32_t u32; typedef int32_t i32;\n#define PGPU_DEV\n"
                "struct P {\n" + src[a:b] + "};\nextern \"C\" int key_order(const u64* a, const u64* b) { return P::key_order(a, b); }\n")
    subprocess.run(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-o", so, cpp], check=True)
    lib = ctypes.CDLL(so)
    lib.key_order.argtypes = [ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
    lib.key_order.restype = ctypes.c_int
    return lib


def pack(row, kinds):
    """the slot's key words (csrc/codegen.cpp: null bits in the first byte(s), then the fields' value bytes back to back, little endian;
    a NULL field's value bytes are zero), with the occupied mark of a tagless slot in the top bit of the last word"""
    nullbyte = 0
    body = b""
    for i, (v, k) in enumerate(zip(row, kinds)):
        if v is None:
            nullbyte |= 1 << i
            body += b"\0" * (4 if k == "i32" else 8)
        elif k == "i64":
            body += struct.pack("<q", v)
        elif k == "i32":
            body += struct.pack("<i", v)
        else:
            body += struct.pack("<d", v + 0.0)   # -0.0 groups with 0.0 (f64_group_bits)
    raw = bytes([nullbyte]) + body
    raw += b"\0" * (-len(raw) % 8)
    words = list(struct.unpack("<%dQ" % (len(raw) // 8), raw))
    words[-1] |= 1 << 63
    return (ctypes.c_uint64 * len(words))(*words)


def sql_order_key(row):
    return tuple((1, 0) if v is None else (0, v) for v in row)   # NULLs last, values by their type's order


@pytest.mark.parametrize("kinds", [("i64",), ("i64", "i32"), ("i32", "f64", "i64")])
def test_key_order_is_the_sql_order_of_the_grouping_fields(kinds):
    types = {"i64": BIGINT, "i32": INTEGER, "f64": DOUBLE}
    inputs = [types[k] for k in kinds] + [DOUBLE]
    plan = plan_text(inputs, step="single", groupByChannels=list(range(len(kinds))),
                     aggregations=[Aggregation("sum", DOUBLE, len(kinds), DOUBLE)], expectedGroups=NO_DIRECTORY)
    lib = key_order_lib(source_of(plan, len(inputs)), "_".join(kinds))
    rng = np.random.default_rng(len(kinds))
    pool = {"i64": [-(1 << 63), -(1 << 40), -2, -1, 0, 1, 2, 255, 256, (1 << 56) - 1, 1 << 56, (1 << 63) - 1],
            "i32": [-(1 << 31), -70000, -1, 0, 1, 255, 256, 65536, (1 << 31) - 1],
            "f64": [float("-inf"), -1e300, -2.5, -1.0, -1e-300, 0.0, 1e-300, 1.0, 2.5, 1e300, float("inf")]}
    rows = set()
    for _ in range(600):
        rows.add(tuple(None if rng.random() < 0.15 else pool[k][int(rng.integers(len(pool[k])))] for k in kinds))
    rows = sorted(rows, key=sql_order_key)
    packed = [pack(r, kinds) for r in rows]
    for i in range(len(rows)):
        assert lib.key_order(packed[i], packed[i]) == 0
        for j in (i + 1, min(i + 7, len(rows) - 1), len(rows) - 1):
            if j <= i or j >= len(rows):
                continue
            assert lib.key_order(packed[i], packed[j]) < 0, (rows[i], rows[j])
            assert lib.key_order(packed[j], packed[i]) > 0, (rows[j], rows[i])


KEY_BYTES = {"date": 4, "integer": 4, "bigint": 8, "boolean": 1, "dec10": 8, "dec30": 16, "varchar5": 6}


@pytest.mark.parametrize("kinds", [("date", "dec10"), ("integer", "varchar5"), ("bigint", "boolean"), ("dec30",), ("boolean", "varchar5", "date")])
def test_key_order_is_a_total_order_that_separates_different_keys(kinds):
    """what the emit kernel relies on for ANY key type: key_order(a, b) == 0 exactly when the key bytes are equal (else two groups
    would merge), it is antisymmetric and transitive (else a key could reappear unnoticed). Value bytes are arbitrary bit patterns
    here (non-NULL fields; the layout is 1 byte of null bits, then the fields back to back)."""
    from presto_b200.spi import BOOLEAN, DATE, createDecimalType, createVarcharType
    types = {"date": DATE, "integer": INTEGER, "bigint": BIGINT, "boolean": BOOLEAN, "dec10": createDecimalType(10, 2),
             "dec30": createDecimalType(30, 2), "varchar5": createVarcharType(5)}
    inputs = [types[k] for k in kinds] + [DOUBLE]
    plan = plan_text(inputs, step="single", groupByChannels=list(range(len(kinds))),
                     aggregations=[Aggregation("sum", DOUBLE, len(kinds), DOUBLE)], expectedGroups=NO_DIRECTORY)
    lib = key_order_lib(source_of(plan, len(inputs)), "total_" + "_".join(kinds))
    rng = np.random.default_rng(7)
    alphabet = np.array([0x00, 0x01, 0x7f, 0x80, 0xff], np.uint8)   # sign bits, carries between bytes

    def random_key():
        raw = bytes([0])
        for k in kinds:
            n = KEY_BYTES[k]
            b = alphabet[rng.integers(0, len(alphabet), n)].tobytes() if k != "boolean" else bytes([int(rng.integers(0, 2))])
            raw += b
        raw += b"\0" * (-len(raw) % 8)
        if len(raw) == 1 + sum(KEY_BYTES[k] for k in kinds):   # the key fills its words: (not a tagless layout; no mark to put)
            return raw, (ctypes.c_uint64 * (len(raw) // 8))(*struct.unpack("<%dQ" % (len(raw) // 8), raw))
        words = list(struct.unpack("<%dQ" % (len(raw) // 8), raw))
        words[-1] |= 1 << 63
        return raw, (ctypes.c_uint64 * len(words))(*words)

    keys = [random_key() for _ in range(300)]
    keys += keys[:40]                                            # equal keys occur
    import functools
    order = sorted(range(len(keys)), key=functools.cmp_to_key(lambda i, j: lib.key_order(keys[i][1], keys[j][1])))
    for a, b in zip(order, order[1:]):                           # sorted by key_order: neighbours never decrease, equal only if the same bytes
        o = lib.key_order(keys[a][1], keys[b][1])
        assert o <= 0 and lib.key_order(keys[b][1], keys[a][1]) == -o
        assert (o == 0) == (keys[a][0] == keys[b][0]), (keys[a][0].hex(), keys[b][0].hex())
    for _ in range(3000):                                        # and any pair agrees with its position in that order
        i, j = (int(x) for x in rng.integers(0, len(order), 2))
        o = lib.key_order(keys[order[i]][1], keys[order[j]][1])
        if keys[order[i]][0] == keys[order[j]][0]:
            assert o == 0
        else:
            assert (o < 0) == (i < j), (keys[order[i]][0].hex(), keys[order[j]][0].hex())
