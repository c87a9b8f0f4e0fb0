"""Host-side mirror of the trino-spi data model the GPU path accepts: Types, Blocks, Page.

Follows core/trino-spi/src/main/java/io/trino/spi/ (reference):
  Page                      Page.java:33-120
  LongArrayBlock            block/LongArrayBlock.java:37-41      (BIGINT, DOUBLE raw bits, short DECIMAL)
  IntArrayBlock             block/IntArrayBlock.java:36-40       (INTEGER, DATE)
  ByteArrayBlock            block/ByteArrayBlock.java:36-40      (BOOLEAN)
  Int128ArrayBlock          block/Int128ArrayBlock.java:35-41    (long DECIMAL, sign-magnitude, low word first)
  VariableWidthBlock        block/VariableWidthBlock.java:38-43  (VARCHAR / VARBINARY: bytes + offsets[n+1])
  DictionaryBlock           block/DictionaryBlock.java:40-48
  RunLengthEncodedBlock     block/RunLengthEncodedBlock.java:45-48
Blocks hold numpy arrays (the Java heap arrays); `valueIsNull` is a uint8/bool array or None.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np

from . import _native as N


@dataclass(frozen=True)
class Type:
    kind: int
    precision: int = 0
    scale: int = 0

    def __str__(self):
        names = {0: "boolean", 1: "bigint", 2: "integer", 3: "date", 4: "double", 7: "varbinary"}
        if self.kind == 5:
            return "(decimal %d %d)" % (self.precision, self.scale)
        if self.kind == 6:
            return "(varchar %d)" % self.precision if self.precision else "varchar"
        return names[self.kind]

    @property
    def is_short_decimal(self):
        return self.kind == 5 and self.precision <= 18

    @property
    def is_long_decimal(self):
        return self.kind == 5 and self.precision > 18

    @property
    def is_var_width(self):
        return self.kind in (6, 7)


BOOLEAN = Type(0)
BIGINT = Type(1)
INTEGER = Type(2)
DATE = Type(3)
DOUBLE = Type(4)
VARBINARY = Type(7)
VARCHAR = Type(6, 0)


def createDecimalType(precision: int, scale: int = 0) -> Type:
    return Type(5, precision, scale)


def createVarcharType(length: int) -> Type:
    return Type(6, length)


SIGN_LONG_MASK = 1 << 63


def encode_int128(values: Sequence[int]) -> np.ndarray:
    """Python ints -> Int128ArrayBlock layout (UnscaledDecimal128Arithmetic.java:30-56)."""
    out = np.zeros((len(values), 2), dtype=np.uint64)
    for i, v in enumerate(values):
        m = -v if v < 0 else v
        if m >> 127:
            raise OverflowError("decimal magnitude needs more than 127 bits")
        out[i, 0] = m & 0xFFFFFFFFFFFFFFFF
        out[i, 1] = (m >> 64) | (SIGN_LONG_MASK if v < 0 else 0)
    return out


def decode_int128(raw: np.ndarray) -> List[int]:
    raw = np.asarray(raw, dtype=np.uint64).reshape(-1, 2)
    out = []
    for lo, hi in raw.tolist():
        m = ((hi & (SIGN_LONG_MASK - 1)) << 64) | lo
        out.append(-m if hi & SIGN_LONG_MASK else m)
    return out


class Block:
    """spi/block/Block.java — only what the operator boundary needs."""
    type: Type
    position_count: int

    def getPositionCount(self):
        return self.position_count

    def mayHaveNull(self):
        return False

    def to_pylist(self) -> list:
        raise NotImplementedError

    # -- marshalling: fills a pgpu_column and returns objects that must stay alive
    def _fill(self, col: N.Column, keep: list):
        raise NotImplementedError


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data


class _FlatBlock(Block):
    dtype = None

    def __init__(self, type: Type, values, valueIsNull=None):
        self.type = type
        self.values = np.ascontiguousarray(values, dtype=self.dtype)
        self.position_count = int(self.values.shape[0])
        if valueIsNull is not None:
            valueIsNull = np.ascontiguousarray(np.asarray(valueIsNull).astype(np.uint8))
            assert valueIsNull.shape[0] == self.position_count
        self.valueIsNull = valueIsNull

    def mayHaveNull(self):
        return self.valueIsNull is not None

    def _fill(self, col, keep):
        col.encoding = N.ENC_PLAIN
        col.type_kind, col.precision, col.scale = self.type.kind, self.type.precision, self.type.scale
        col.position_count = self.position_count
        col.values = _ptr(self.values)
        col.nulls = _ptr(self.valueIsNull)
        keep.append(self)

    def _py(self, i):
        return self.values[i].item()

    def to_pylist(self):
        nulls = self.valueIsNull
        return [None if (nulls is not None and nulls[i]) else self._py(i) for i in range(self.position_count)]


class LongArrayBlock(_FlatBlock):
    dtype = np.int64

    def __init__(self, type: Type, values, valueIsNull=None):
        if type.kind == 4:
            values = np.ascontiguousarray(values, dtype=np.float64).view(np.int64)
        super().__init__(type, values, valueIsNull)

    def _py(self, i):
        if self.type.kind == 4:
            return self.values.view(np.float64)[i].item()
        return int(self.values[i])


class IntArrayBlock(_FlatBlock):
    dtype = np.int32


class ByteArrayBlock(_FlatBlock):
    dtype = np.int8

    def _py(self, i):
        return bool(self.values[i]) if self.type.kind == 0 else int(self.values[i])


class Int128ArrayBlock(_FlatBlock):
    dtype = np.uint64

    def __init__(self, type: Type, values, valueIsNull=None):
        """values: (n, 2) uint64 in block layout, or a sequence of Python ints."""
        if not isinstance(values, np.ndarray):
            values = encode_int128([0 if v is None else v for v in values])
        super().__init__(type, np.asarray(values, dtype=np.uint64).reshape(-1, 2), valueIsNull)

    def _py(self, i):
        return decode_int128(self.values[i])[0]


class VariableWidthBlock(Block):
    def __init__(self, type: Type, data: bytes | np.ndarray, offsets, valueIsNull=None):
        self.type = type
        self.data = np.frombuffer(bytes(data), dtype=np.uint8).copy() if not isinstance(data, np.ndarray) else np.ascontiguousarray(data, dtype=np.uint8)
        if self.data.shape[0] == 0:
            self.data = np.zeros(1, dtype=np.uint8)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int32)
        self.position_count = int(self.offsets.shape[0]) - 1
        if valueIsNull is not None:
            valueIsNull = np.ascontiguousarray(np.asarray(valueIsNull).astype(np.uint8))
        self.valueIsNull = valueIsNull

    @staticmethod
    def from_values(type: Type, values: Sequence[Optional[bytes | str]]):
        enc = [b"" if v is None else (v.encode() if isinstance(v, str) else bytes(v)) for v in values]
        offsets = np.zeros(len(enc) + 1, dtype=np.int32)
        np.cumsum([len(e) for e in enc], out=offsets[1:])
        nulls = np.array([v is None for v in values], dtype=np.uint8)
        return VariableWidthBlock(type, b"".join(enc), offsets, nulls if nulls.any() else None)

    def mayHaveNull(self):
        return self.valueIsNull is not None

    def _fill(self, col, keep):
        col.encoding = N.ENC_PLAIN
        col.type_kind, col.precision, col.scale = self.type.kind, self.type.precision, self.type.scale
        col.position_count = self.position_count
        col.values = _ptr(self.data)
        col.offsets = _ptr(self.offsets)
        col.nulls = _ptr(self.valueIsNull)
        keep.append(self)

    def to_pylist(self):
        out = []
        raw = self.data.tobytes()
        for i in range(self.position_count):
            if self.valueIsNull is not None and self.valueIsNull[i]:
                out.append(None)
            else:
                b = raw[self.offsets[i]:self.offsets[i + 1]]
                out.append(b.decode() if self.type.kind == 6 else b)
        return out


class DictionaryBlock(Block):
    def __init__(self, dictionary: Block, ids):
        assert not isinstance(dictionary, (DictionaryBlock, RunLengthEncodedBlock))
        self.type = dictionary.type
        self.dictionary = dictionary
        self.ids = np.ascontiguousarray(ids, dtype=np.int32)
        self.position_count = int(self.ids.shape[0])

    def mayHaveNull(self):
        return self.dictionary.mayHaveNull()

    def _fill(self, col, keep):
        d = N.Column()
        self.dictionary._fill(d, keep)
        keep.append(d)
        col.encoding = N.ENC_DICTIONARY
        col.type_kind, col.precision, col.scale = self.type.kind, self.type.precision, self.type.scale
        col.position_count = self.position_count
        col.ids = _ptr(self.ids)
        col.dictionary = ctypes.pointer(d)
        keep.append(self)

    def to_pylist(self):
        d = self.dictionary.to_pylist()
        return [d[i] for i in self.ids.tolist()]


class RunLengthEncodedBlock(Block):
    def __init__(self, value: Block, positionCount: int):
        assert value.position_count == 1
        self.type = value.type
        self.value = value
        self.position_count = int(positionCount)

    def mayHaveNull(self):
        return self.value.mayHaveNull()

    def _fill(self, col, keep):
        d = N.Column()
        self.value._fill(d, keep)
        keep.append(d)
        col.encoding = N.ENC_RLE
        col.type_kind, col.precision, col.scale = self.type.kind, self.type.precision, self.type.scale
        col.position_count = self.position_count
        col.dictionary = ctypes.pointer(d)
        keep.append(self)

    def to_pylist(self):
        return self.value.to_pylist() * self.position_count


def block_of(type: Type, values: Sequence) -> Block:
    """Build a flat block from Python values (None = NULL) — the RowPagesBuilder of the tests
    (core/trino-main/src/test/java/io/trino/RowPagesBuilder.java)."""
    nulls = np.array([v is None for v in values], dtype=np.uint8)
    nn = nulls if nulls.any() else None
    if type.is_var_width:
        return VariableWidthBlock.from_values(type, values)
    if type.kind == 0:
        return ByteArrayBlock(type, [0 if v is None else int(bool(v)) for v in values], nn)
    if type.kind in (2, 3):
        return IntArrayBlock(type, [0 if v is None else v for v in values], nn)
    if type.kind == 4:
        return LongArrayBlock(type, np.array([0.0 if v is None else v for v in values], dtype=np.float64), nn)
    if type.is_long_decimal:
        return Int128ArrayBlock(type, [0 if v is None else v for v in values], nn)
    return LongArrayBlock(type, [0 if v is None else v for v in values], nn)


class Page:
    """spi/Page.java:33-120"""

    def __init__(self, *blocks: Block, positionCount: Optional[int] = None):
        self.blocks = list(blocks)
        if positionCount is None:
            positionCount = blocks[0].position_count if blocks else 0
        self.position_count = positionCount
        for b in self.blocks:
            assert b.position_count == positionCount, "block position counts differ"

    def getPositionCount(self):
        return self.position_count

    def getChannelCount(self):
        return len(self.blocks)

    def getBlock(self, channel):
        return self.blocks[channel]

    def to_rows(self) -> List[tuple]:
        cols = [b.to_pylist() for b in self.blocks]
        return [tuple(c[i] for c in cols) for i in range(self.position_count)]


def block_from_column(col: N.Column) -> Block:
    """Copy a library-owned output block (valid until the next get_output) into a Python-owned Block."""
    n = col.position_count
    t = Type(col.type_kind, col.precision, col.scale)
    nulls = None
    if col.nulls and n:
        nb = np.ctypeslib.as_array(ctypes.cast(col.nulls, ctypes.POINTER(ctypes.c_uint8)), shape=(n,)).copy()
        nulls = nb if nb.any() else None

    def arr(ctype, dtype, count):
        if count == 0 or not col.values:
            return np.zeros(count, dtype=dtype)
        return np.ctypeslib.as_array(ctypes.cast(col.values, ctypes.POINTER(ctype)), shape=(count,)).astype(dtype, copy=True)

    if t.is_var_width:
        offs = np.ctypeslib.as_array(ctypes.cast(col.offsets, ctypes.POINTER(ctypes.c_int32)), shape=(n + 1,)).copy() if n else np.zeros(1, np.int32)
        data = arr(ctypes.c_uint8, np.uint8, int(offs[-1]))
        return VariableWidthBlock(t, data, offs, nulls)
    if t.kind == 0:
        return ByteArrayBlock(t, arr(ctypes.c_int8, np.int8, n), nulls)
    if t.kind in (2, 3):
        return IntArrayBlock(t, arr(ctypes.c_int32, np.int32, n), nulls)
    if t.is_long_decimal:
        return Int128ArrayBlock(t, arr(ctypes.c_uint64, np.uint64, 2 * n).reshape(-1, 2), nulls)
    if t.kind == 4:
        return LongArrayBlock(t, arr(ctypes.c_double, np.float64, n), nulls)
    return LongArrayBlock(t, arr(ctypes.c_int64, np.int64, n), nulls)
