"""TEST INFRASTRUCTURE — CPU restatement of the reference's page wire format (writer and reader), used only by tests/ to
produce and check serialized pages. The product never imports this.

Follows (electrum/presto = Trino 359):
  * main/execution/buffer/PagesSerdeUtil.java:45-62   writeRawPage / readRawPage: int32 channel count, then every block
  * main/execution/buffer/PagesSerdeUtil.java:64-72   writeSerializedPage: positionCount, markers, uncompressed size, size, payload
  * main/metadata/InternalBlockEncodingSerde.java:43-80  length-prefixed encoding name, then the encoding's body
  * spi/block/EncoderUtil.java:33-70                  null flags as a bit stream, most significant bit first
  * spi/block/LongArrayBlockEncoding.java:37-61 (Int / Byte / Int128 alike): positionCount, nulls, then either all values
    or (when the block may have nulls) nonNullCount + the non-null values only
  * spi/block/VariableWidthBlockEncoding.java:37-58   positionCount, END offsets (no leading 0), nulls, total length, bytes
  * spi/block/DictionaryBlockEncoding.java:36-58      positionCount, dictionary block, int32 ids, 3 x int64 dictionary id
  * spi/block/RunLengthBlockEncoding.java:31-40       positionCount, value block
All integers little endian (airlift Slice). Pinned by the serialized-size known answers of
main/test/execution/buffer/TestPagesSerde.java:64-108 (tests/test_pages_serde.py) plus round trips.

Blocks are read through duck typing (`values`, `valueIsNull`, `offsets`, `data`, `ids`, `dictionary`, `value`,
`position_count`); the reader returns plain dicts so that this file depends on numpy only."""
import struct

import numpy as np

FLAT_NAMES = {1: b"BYTE_ARRAY", 4: b"INT_ARRAY", 8: b"LONG_ARRAY", 16: b"INT128_ARRAY"}
FLAT_WIDTH = {v: k for k, v in FLAT_NAMES.items()}


def _name(b: bytes) -> bytes:
    return struct.pack("<i", len(b)) + b


def encode_nulls_as_bits(nulls, n: int) -> bytes:
    """EncoderUtil.encodeNullsAsBits: mayHaveNull byte, then ceil(n / 8) bytes, position p -> bit 7 - p % 8 of byte p / 8"""
    if nulls is None:
        return b"\x00"
    bits = np.packbits(np.asarray(nulls[:n], dtype=np.uint8) != 0, bitorder="big")
    return b"\x01" + bits.tobytes()


def _write_flat(block) -> bytes:
    values = np.ascontiguousarray(block.values)
    n = block.position_count
    width = values.dtype.itemsize * (2 if values.ndim == 2 else 1)   # Int128ArrayBlock stores (n, 2) 64-bit words
    raw = values.view(np.uint8).reshape(n, width)
    nulls = getattr(block, "valueIsNull", None)
    out = [_name(FLAT_NAMES[width]), struct.pack("<i", n), encode_nulls_as_bits(nulls, n)]
    if nulls is None:
        out.append(raw.tobytes())
    else:
        keep = np.asarray(nulls[:n]) == 0
        out.append(struct.pack("<i", int(keep.sum())))
        out.append(raw[keep].tobytes())
    return b"".join(out)


def write_block(block) -> bytes:
    if hasattr(block, "ids"):          # DictionaryBlock
        n = block.position_count
        return b"".join([_name(b"DICTIONARY"), struct.pack("<i", n), write_block(block.dictionary),
                         np.ascontiguousarray(block.ids, dtype="<i4").tobytes(), struct.pack("<qqq", 0x1234, 0x5678, 1)])
    if hasattr(block, "value"):        # RunLengthEncodedBlock
        return b"".join([_name(b"RLE"), struct.pack("<i", block.position_count), write_block(block.value)])
    if hasattr(block, "offsets"):      # VariableWidthBlock
        n = block.position_count
        offs = np.asarray(block.offsets, dtype=np.int64)
        ends = (offs[1:n + 1] - offs[0]).astype("<i4")
        total = int(ends[-1]) if n else 0
        data = np.asarray(block.data, dtype=np.uint8)[int(offs[0]):int(offs[0]) + total]
        return b"".join([_name(b"VARIABLE_WIDTH"), struct.pack("<i", n), ends.tobytes(),
                         encode_nulls_as_bits(getattr(block, "valueIsNull", None), n), struct.pack("<i", total), data.tobytes()])
    return _write_flat(block)


def write_raw_page(blocks) -> bytes:
    """PagesSerdeUtil.writeRawPage"""
    return struct.pack("<i", len(blocks)) + b"".join(write_block(b) for b in blocks)


def write_serialized_page(blocks, position_count: int) -> bytes:
    """PagesSerdeUtil.writeSerializedPage of an uncompressed, unencrypted page (markers = 0)"""
    payload = write_raw_page(blocks)
    return struct.pack("<ibii", position_count, 0, len(payload), len(payload)) + payload


class _In:
    def __init__(self, data: bytes):
        self.b, self.p = data, 0

    def i32(self) -> int:
        v = struct.unpack_from("<i", self.b, self.p)[0]
        self.p += 4
        return v

    def take(self, n: int) -> bytes:
        v = self.b[self.p:self.p + n]
        assert len(v) == n, "truncated page"
        self.p += n
        return v


def decode_null_bits(inp: _In, n: int):
    """EncoderUtil.retrieveNullBits + decodeNullBits -> uint8[n] or None"""
    if inp.take(1) == b"\x00":
        return None
    packed = np.frombuffer(inp.take((n + 7) // 8), dtype=np.uint8)
    return np.unpackbits(packed, bitorder="big")[:n].astype(np.uint8)


def read_block(inp: _In) -> dict:
    name = inp.take(inp.i32())
    n = inp.i32()
    if name in FLAT_WIDTH:
        w = FLAT_WIDTH[name]
        nulls = decode_null_bits(inp, n)
        if nulls is None:
            raw = np.frombuffer(inp.take(n * w), dtype=np.uint8).reshape(n, w)
        else:
            nn = inp.i32()
            assert nn == int((nulls == 0).sum())
            compact = np.frombuffer(inp.take(nn * w), dtype=np.uint8).reshape(nn, w)
            raw = np.zeros((n, w), dtype=np.uint8)
            raw[nulls == 0] = compact
        return {"encoding": name.decode(), "positionCount": n, "width": w, "valueIsNull": nulls, "raw": raw}
    if name == b"VARIABLE_WIDTH":
        offsets = np.zeros(n + 1, dtype=np.int32)
        offsets[1:] = np.frombuffer(inp.take(4 * n), dtype="<i4")
        nulls = decode_null_bits(inp, n)
        total = inp.i32()
        return {"encoding": "VARIABLE_WIDTH", "positionCount": n, "offsets": offsets, "valueIsNull": nulls,
                "data": np.frombuffer(inp.take(total), dtype=np.uint8)}
    if name == b"DICTIONARY":
        d = read_block(inp)
        ids = np.frombuffer(inp.take(4 * n), dtype="<i4").astype(np.int32)
        inp.take(24)
        return {"encoding": "DICTIONARY", "positionCount": n, "dictionary": d, "ids": ids}
    if name == b"RLE":
        return {"encoding": "RLE", "positionCount": n, "value": read_block(inp)}
    raise ValueError("unsupported block encoding %r" % name)


def read_raw_page(data: bytes) -> list:
    inp = _In(data)
    blocks = [read_block(inp) for _ in range(inp.i32())]
    assert inp.p == len(data), "trailing bytes"
    return blocks


def block_values(b: dict) -> list:
    """decoded block -> Python values of the raw storage (ints for flat blocks by width, bytes for var-width, None = NULL)"""
    n = b["positionCount"]
    if b["encoding"] == "DICTIONARY":
        d = block_values(b["dictionary"])
        return [d[i] for i in b["ids"].tolist()]
    if b["encoding"] == "RLE":
        return block_values(b["value"]) * n
    nulls = b["valueIsNull"]
    if b["encoding"] == "VARIABLE_WIDTH":
        raw = b["data"].tobytes()
        vals = [raw[b["offsets"][i]:b["offsets"][i + 1]] for i in range(n)]
    elif b["width"] == 16:
        vals = [bytes(r) for r in b["raw"]]
    else:
        vals = b["raw"].copy().view({1: np.int8, 4: "<i4", 8: "<i8"}[b["width"]]).reshape(n).tolist()
    return [None if nulls is not None and nulls[i] else vals[i] for i in range(n)]
