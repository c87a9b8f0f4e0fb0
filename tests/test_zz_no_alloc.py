"""pgpu_expand_nulls_static — the allocation-free entry point (scratch in a file-scope __device__ array of the library,
no counterpart in the reference) against the existing path that runs the same kernels with
allocated scratch: pgpu_add_serialized_page decoding a nullable fixed-width block, observed through an identity filter + project
pipeline and pgpu_get_output. Byte / integer work: bit-exact. Two calls in a row (the second reuses the array with the first
call's contents still in it), sizes around the 1024-position block boundaries, every value width.
Collected last on purpose (file name)."""
import ctypes
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import pages_serde as S  # noqa: E402

from presto_b200 import _native as N  # noqa: E402
from presto_b200.operators import GpuPipelineOperator, plan_text  # noqa: E402
from presto_b200.relational import field  # noqa: E402
from presto_b200.spi import (BIGINT, BOOLEAN, DOUBLE, INTEGER, ByteArrayBlock, Int128ArrayBlock, IntArrayBlock, LongArrayBlock, Page,  # noqa: E402
                             createDecimalType)

DEC30 = createDecimalType(30, 2)


class DeviceBuf:
    """test plumbing: a device buffer from pgpu_device_alloc, copies through libprestogpu_bench.so"""

    def __init__(self, nbytes):
        self.nbytes = max(int(nbytes), 16)
        self.ptr = N.lib().pgpu_device_alloc(0, self.nbytes)
        assert self.ptr

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes <= self.nbytes
        if arr.nbytes:
            assert N.lib().pgpu_memcpy_h2d(0, self.ptr, arr.ctypes.data, arr.nbytes) == 0
        return self

    def download(self, dtype, count):
        out = np.empty(count, dtype=dtype)
        if out.nbytes:
            assert N.lib().pgpu_memcpy_d2h(0, out.ctypes.data, self.ptr, out.nbytes) == 0   # synchronous: orders after the default stream
        return out

    def free(self):
        N.lib().pgpu_device_free(0, self.ptr)


def make_block(kind, n, seed):
    rng = np.random.default_rng(seed)
    nulls = (rng.random(n) < 0.35).astype(np.uint8)
    if n > 3:
        nulls[0], nulls[n - 1] = 1, 0
    if kind == "bigint":
        return BIGINT, LongArrayBlock(BIGINT, rng.integers(-2**62, 2**62, n), nulls), 8
    if kind == "double":
        return DOUBLE, LongArrayBlock(DOUBLE, rng.random(n) * 1e6 - 5e5, nulls), 8
    if kind == "integer":
        return INTEGER, IntArrayBlock(INTEGER, rng.integers(-2**31, 2**31 - 1, n), nulls), 4
    if kind == "boolean":
        return BOOLEAN, ByteArrayBlock(BOOLEAN, rng.integers(0, 2, n), nulls), 1
    return DEC30, Int128ArrayBlock(DEC30, [int(x) * 10**11 + 7 for x in rng.integers(-10**9, 10**9, n)], nulls), 16


def existing_path(type_, block, n):
    """the same block through pgpu_add_serialized_page (expandNulls with allocated scratch) -> identity projection -> pgpu_get_output"""
    op = GpuPipelineOperator(plan_text([type_], None, [field(0, type_)]), 0)
    try:
        op.addSerializedInput(S.write_raw_page([block]), n)
        out = op.getOutput()
        assert out is not None and out.position_count == n
        b = out.blocks[0]
        nulls = b.valueIsNull if b.valueIsNull is not None else np.zeros(n, dtype=np.uint8)
        return np.ascontiguousarray(b.values).view(np.uint8).reshape(n, -1), np.asarray(nulls, dtype=np.uint8)
    finally:
        op.close()


def static_path(block, n, width):
    """pgpu_expand_nulls_static on the pieces the wire carries: packed null bits + the non-null values only"""
    L = N.lib()
    nulls = np.asarray(block.valueIsNull, dtype=np.uint8)
    raw = np.ascontiguousarray(block.values).view(np.uint8).reshape(n, width)
    bits = np.packbits(nulls != 0, bitorder="big")                     # EncoderUtil.encodeNullsAsBits: most significant bit first
    compact = np.ascontiguousarray(raw[nulls == 0])
    dbits, dcompact = DeviceBuf(bits.nbytes).upload(bits), DeviceBuf(compact.nbytes).upload(compact)
    dvalues, dnulls = DeviceBuf(n * width), DeviceBuf(n)
    dvalues.upload(np.full(n * width, 0xA5, dtype=np.uint8))           # stale bytes the call must overwrite everywhere
    dnulls.upload(np.full(n, 0xA5, dtype=np.uint8))
    try:
        st = L.pgpu_expand_nulls_static(0, None, dbits.ptr, n, dcompact.ptr, width, dvalues.ptr, dnulls.ptr)
        assert st == N.OK, L.pgpu_last_global_error().decode()
        return dvalues.download(np.uint8, n * width).reshape(n, width), dnulls.download(np.uint8, n)
    finally:
        for b in (dbits, dcompact, dvalues, dnulls):
            b.free()


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["bigint", "double", "integer", "boolean", "decimal30"])
def test_static_scratch_equals_existing_path_twice_in_a_row(kind):
    # a large call first, then a small one: the second finds the first call's block counts in the shared array and must not read them
    for call, n in enumerate((70_001, 1025, 1024, 1023, 9, 1)):
        type_, block, width = make_block(kind, n, seed=100 * call + n)
        want_values, want_nulls = existing_path(type_, block, n)
        got_values, got_nulls = static_path(block, n, width)
        assert np.array_equal(got_nulls, want_nulls), "call %d (n = %d): null flags differ" % (call, n)
        live = want_nulls == 0
        assert np.array_equal(got_values[live], want_values[live]), "call %d (n = %d): values differ" % (call, n)
        assert not got_values[~live].any(), "call %d (n = %d): NULL positions must be zero" % (call, n)
        # and both against the block itself
        assert np.array_equal(got_nulls, np.asarray(block.valueIsNull, dtype=np.uint8))
        assert np.array_equal(got_values[live], np.ascontiguousarray(block.values).view(np.uint8).reshape(n, width)[live])


@pytest.mark.gpu
def test_static_scratch_limits_are_errors():
    L = N.lib()
    d = DeviceBuf(64)
    try:
        too_many = 16_777_216 + 1                                       # PGPU_EXPAND_STATIC_MAX_POSITIONS + 1
        assert L.pgpu_expand_nulls_static(0, None, d.ptr, too_many, d.ptr, 8, d.ptr, d.ptr) == N.INVALID_ARGUMENT
        msg = L.pgpu_last_global_error().decode()
        assert "exceeds" in msg and "16777216" in msg
        assert L.pgpu_expand_nulls_static(0, None, d.ptr, 10, d.ptr, 3, d.ptr, d.ptr) == N.INVALID_ARGUMENT       # width
        assert L.pgpu_expand_nulls_static(0, None, d.ptr, -1, d.ptr, 8, d.ptr, d.ptr) == N.INVALID_ARGUMENT      # negative count
        assert L.pgpu_expand_nulls_static(0, None, None, 10, d.ptr, 8, d.ptr, d.ptr) == N.INVALID_ARGUMENT       # null pointer
        assert L.pgpu_expand_nulls_static(0, None, d.ptr, 0, d.ptr, 8, d.ptr, d.ptr) == N.OK                     # an empty block is fine
    finally:
        d.free()
