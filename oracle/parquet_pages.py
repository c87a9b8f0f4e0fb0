"""CPU oracle for Parquet column chunks -> Blocks (TEST INFRASTRUCTURE: only tests/ may import it; the product path is
presto_b200/csrc/parquet_host.h + parquet.cu and never calls into this file).

Restates what the reference does between a column chunk's bytes and the Block it hands to the scan operator:

  lib/trino-parquet/src/main/java/io/trino/parquet/reader/ParquetColumnChunk.java:60-135   page headers, dictionary / data pages
  lib/trino-parquet/src/main/java/io/trino/parquet/reader/PageReader.java:55-109            (de)compression, page iteration
  lib/trino-parquet/src/main/java/io/trino/parquet/reader/PrimitiveColumnReader.java:148-317  levels + values -> BlockBuilder
  lib/trino-parquet/src/main/java/io/trino/parquet/ParquetEncoding.java:58-234              PLAIN / RLE / *_DICTIONARY readers
  lib/trino-parquet/src/main/java/io/trino/parquet/dictionary/{Long,Integer,Double,Binary}Dictionary.java
  lib/trino-parquet/src/main/java/io/trino/parquet/reader/{Long,Int,Double,Boolean,Binary}ColumnReader.java

The arithmetic itself lives in un-vendored third-party code: the org.apache.parquet classes (ValuesReader, RunLengthBitPacking-
HybridDecoder, PlainValuesReader, BytePacker, the Thrift compact PageHeader reader) shaded into io.trino.hive:hive-apache
3.1.2-12 (/root/reference/pom.xml:546-547, lib/trino-parquet/pom.xml; imported at PrimitiveColumnReader.java:31-36). Their published format (parquet-format Encodings.md: PLAIN = little-endian values,
BYTE_ARRAY = 4-byte length + bytes, BOOLEAN = bits LSB first; RLE/bit-packed hybrid = varint header, header & 1 ? (header >> 1)
groups of 8 values bit-packed LSB first : (header >> 1) repeats of one ceil(width / 8)-byte little-endian value; data page v1 =
[4-byte length + hybrid repetition levels][4-byte length + hybrid definition levels][values], v2 = levels without the length
prefix, sizes in the header; dictionary ids = 1 byte bit width + hybrid) is restated here and pinned against an independent
implementation, pyarrow's C++ reader, on files pyarrow's writer produced (tests/test_parquet_oracle.py, tests/golden/*.parquet).

Subset (everything else raises Unsupported, as the product rejects it): flat columns (max repetition level 0, max definition
level <= 1), UNCOMPRESSED codec, data pages v1 and v2, encodings PLAIN, PLAIN_DICTIONARY, RLE_DICTIONARY (+ RLE for levels and BOOLEAN values),
physical types BOOLEAN, INT32, INT64, DOUBLE, BYTE_ARRAY (dictionary-encoded chunks), FIXED_LEN_BYTE_ARRAY of 1..16 bytes
as a DECIMAL's unscaled value.
"""
import struct

import numpy as np

BOOLEAN, INT32, INT64, INT96, FLOAT, DOUBLE, BYTE_ARRAY, FIXED_LEN_BYTE_ARRAY = range(8)   # parquet.thrift Type
ENC_PLAIN, ENC_PLAIN_DICTIONARY, ENC_RLE, ENC_BIT_PACKED, ENC_RLE_DICTIONARY = 0, 2, 3, 4, 8   # parquet.thrift Encoding
PAGE_DATA, PAGE_INDEX, PAGE_DICTIONARY, PAGE_DATA_V2 = range(4)                            # parquet.thrift PageType


class Unsupported(Exception):
    pass


class Corrupt(Exception):
    pass


# ------------------------------------------------------------------------------------------- Thrift compact protocol
class Compact:
    """Just enough of the Thrift compact protocol to read parquet.thrift's PageHeader and skip what is not needed."""
    def __init__(self, buf, pos):
        self.b, self.p = buf, pos

    def byte(self):
        if self.p >= len(self.b):
            raise Corrupt("truncated page header")
        v = self.b[self.p]
        self.p += 1
        return v

    def varint(self):
        v, s = 0, 0
        while True:
            c = self.byte()
            v |= (c & 0x7F) << s
            if not c & 0x80:
                return v
            s += 7
            if s > 63:
                raise Corrupt("varint too long")

    def zigzag(self):
        v = self.varint()
        return (v >> 1) ^ -(v & 1)

    def skip(self, t):
        if t in (1, 2):          # BOOLEAN_TRUE / BOOLEAN_FALSE: value is in the field header
            return
        if t == 3:
            self.byte()
        elif t in (4, 5, 6):
            self.varint()
        elif t == 7:
            self.p += 8
        elif t == 8:
            n = self.varint()
            self.p += n
        elif t in (9, 10):       # list / set
            h = self.byte()
            n = h >> 4
            if n == 15:
                n = self.varint()
            for _ in range(n):
                self.skip_element(h & 15)
        elif t == 11:            # map
            n = self.varint()
            if n:
                kv = self.byte()
                for _ in range(n):
                    self.skip_element(kv >> 4)
                    self.skip_element(kv & 15)
        elif t == 12:
            self.struct(lambda fid, ft: self.skip(ft))
        else:
            raise Corrupt("unknown thrift type %d" % t)

    def skip_element(self, t):
        if t in (1, 2):          # booleans inside collections take one byte
            self.byte()
        else:
            self.skip(t)

    def struct(self, on_field):
        last = 0
        while True:
            h = self.byte()
            if h == 0:
                return
            t, delta = h & 15, h >> 4
            fid = last + delta if delta else self.zigzag()
            last = fid
            on_field(fid, t)


def read_page_header(buf, pos):
    """-> (header dict, position of the page body). parquet.thrift: PageHeader{1 type, 2 uncompressed_page_size,
    3 compressed_page_size, 4 crc, 5 data_page_header, 6 index_page_header, 7 dictionary_page_header, 8 data_page_header_v2};
    DataPageHeader{1 num_values, 2 encoding, 3 definition_level_encoding, 4 repetition_level_encoding, 5 statistics};
    DictionaryPageHeader{1 num_values, 2 encoding, 3 is_sorted}; DataPageHeaderV2{1 num_values, 2 num_nulls, 3 num_rows,
    4 encoding, 5 definition_levels_byte_length, 6 repetition_levels_byte_length, 7 is_compressed, 8 statistics}"""
    c = Compact(buf, pos)
    h = {}

    def sub(prefix, names):
        def on(fid, t):
            if fid in names and t in (4, 5, 6):
                h[prefix + names[fid]] = c.zigzag()
            elif fid in names and t in (1, 2):
                h[prefix + names[fid]] = t == 1
            else:
                c.skip(t)
        c.struct(on)

    def top(fid, t):
        if fid in (1, 2, 3) and t == 5:
            h[{1: "type", 2: "uncompressed_size", 3: "compressed_size"}[fid]] = c.zigzag()
        elif fid == 5 and t == 12:
            sub("", {1: "num_values", 2: "encoding", 3: "def_encoding", 4: "rep_encoding"})
        elif fid == 7 and t == 12:
            sub("", {1: "num_values", 2: "encoding"})
        elif fid == 8 and t == 12:
            sub("", {1: "num_values", 2: "num_nulls", 3: "num_rows", 4: "encoding", 5: "def_bytes", 6: "rep_bytes", 7: "is_compressed"})
        else:
            c.skip(t)
    c.struct(top)
    for k in ("type", "uncompressed_size", "compressed_size"):
        if k not in h:
            raise Corrupt("page header without " + k)
    return h, c.p


# ------------------------------------------------------------------------------------------- RLE / bit-packed hybrid
def hybrid_decode(buf, pos, end, width, count):
    """-> uint32[count] (RunLengthBitPackingHybridDecoder). Values past `count` in the last bit-packed group are padding."""
    out = np.empty(count, np.uint32)
    n = 0
    nbytes = (width + 7) // 8
    while n < count:
        c = Compact(buf, pos)
        if pos >= end:
            raise Corrupt("hybrid stream ends after %d of %d values" % (n, count))
        h = c.varint()
        pos = c.p
        if h & 1:
            groups = h >> 1
            take = min(groups * 8, count - n)
            bits = np.unpackbits(np.frombuffer(buf, np.uint8, min(groups * width, end - pos), pos), bitorder="little")
            if len(bits) < take * width:
                raise Corrupt("bit-packed run exceeds the stream")
            vals = bits[:take * width].reshape(take, width).astype(np.uint32) if width else np.zeros((take, 0), np.uint32)
            out[n:n + take] = (vals << np.arange(width, dtype=np.uint32)).sum(axis=1, dtype=np.uint32) if width else 0
            pos += groups * width
            n += take
        else:
            run = h >> 1
            if run == 0:
                raise Corrupt("empty RLE run")
            if pos + nbytes > end:
                raise Corrupt("RLE run exceeds the stream")
            v = int.from_bytes(buf[pos:pos + nbytes], "little")
            pos += nbytes
            take = min(run, count - n)
            out[n:n + take] = v
            n += take
    return out


# ------------------------------------------------------------------------------------------- values
_FIXED = {INT32: ("<i4", 4), INT64: ("<i8", 8), DOUBLE: ("<f8", 8)}


def plain_values(buf, pos, end, ptype, count, type_length=0):
    """PLAIN-encoded values -> numpy array (fixed width), list of bytes (BYTE_ARRAY), or numpy object array of Python ints
    (FIXED_LEN_BYTE_ARRAY read as a DECIMAL's unscaled value: big-endian two's complement, ParquetTypeUtils.getShortDecimalValue
    :226-240 / new BigInteger(bytes) in LongDecimalColumnReader.java:58)"""
    if ptype == FIXED_LEN_BYTE_ARRAY:
        if not 1 <= type_length <= 16:
            raise Unsupported("FIXED_LEN_BYTE_ARRAY(%d) (only DECIMALs of up to 16 bytes)" % type_length)
        if pos + count * type_length > end:
            raise Corrupt("PLAIN values exceed the page")
        out = np.empty(count, object)
        for i in range(count):
            out[i] = int.from_bytes(buf[pos + i * type_length:pos + (i + 1) * type_length], "big", signed=True)
        return out
    if ptype in _FIXED:
        dt, w = _FIXED[ptype]
        if pos + count * w > end:
            raise Corrupt("PLAIN values exceed the page")
        return np.frombuffer(buf, dt, count, pos).copy()
    if ptype == BOOLEAN:
        if pos + (count + 7) // 8 > end:
            raise Corrupt("PLAIN booleans exceed the page")
        return np.unpackbits(np.frombuffer(buf, np.uint8, (count + 7) // 8, pos), bitorder="little")[:count].astype(np.uint8)
    if ptype == BYTE_ARRAY:
        out = []
        for _ in range(count):
            if pos + 4 > end:
                raise Corrupt("BYTE_ARRAY length exceeds the page")
            (n,) = struct.unpack_from("<i", buf, pos)
            pos += 4
            if n < 0 or pos + n > end:
                raise Corrupt("BYTE_ARRAY value exceeds the page")
            out.append(bytes(buf[pos:pos + n]))
            pos += n
        return out
    raise Unsupported("parquet physical type %d" % ptype)


CODEC_NAMES = {1: "snappy", 2: "gzip", 6: "zstd", 7: "lz4_raw"}   # parquet.thrift CompressionCodec -> pyarrow codec name


def _inflate(codec, raw, size):
    """one compressed page body -> bytes, with pyarrow's codec (Arrow's C++ implementation: independent of the product's decoders)"""
    import pyarrow as pa
    if codec not in CODEC_NAMES:
        raise Unsupported("compression codec %d" % codec)
    if size == 0:
        return b""
    return pa.Codec(CODEC_NAMES[codec]).decompress(bytes(raw), decompressed_size=size).to_pybytes()


def decode_chunk(buf, ptype, max_def, num_values, type_length=0, codec=0):
    """One column chunk (bytes from its first page header to its end) -> (values, nulls uint8[num_values], info).
    values: numpy array with 0 at NULL positions (fixed width) / list with None (BYTE_ARRAY).
    info: {"pages": [...], "all_dictionary": bool, "dictionary": array/list or None, "ids": int32[n] with -1 at NULLs (when
    all data pages are dictionary encoded, the form the product turns into a DictionaryBlock)}"""
    if max_def not in (0, 1):
        raise Unsupported("nested columns (max definition level %d)" % max_def)
    buf = bytes(buf)
    pos, done = 0, 0
    dictionary = None
    nulls = np.zeros(num_values, np.uint8)
    if ptype == FIXED_LEN_BYTE_ARRAY:
        values = np.zeros(num_values, object)
    else:
        values = [None] * num_values if ptype == BYTE_ARRAY else np.zeros(num_values, {BOOLEAN: np.uint8, INT32: np.int32, INT64: np.int64, DOUBLE: np.float64}.get(ptype, np.int64))
    ids_all = np.full(num_values, -1, np.int32)
    pages, all_dict = [], True
    chunk = buf
    while done < num_values:
        buf = chunk   # (a compressed page is decoded out of its own uncompressed copy, see below)
        if pos >= len(buf):
            raise Corrupt("column chunk ends after %d of %d values" % (done, num_values))
        h, body = read_page_header(buf, pos)
        end = body + h["compressed_size"]
        if end > len(buf):
            raise Corrupt("page exceeds the column chunk")
        pos = end
        if codec:
            # PageReader.readPage (lib/trino-parquet/.../reader/PageReader.java:63-99): v1 and dictionary pages are compressed whole,
            # v2 pages keep their levels outside the compressed part. The page is decoded out of a private uncompressed copy.
            lv = (h.get("def_bytes", 0) + h.get("rep_bytes", 0)) if h["type"] == PAGE_DATA_V2 else 0
            packed = h.get("is_compressed", True) if h["type"] == PAGE_DATA_V2 else True
            raw = bytes(buf[body:end])
            page = raw[:lv] + (_inflate(codec, raw[lv:], h["uncompressed_size"] - lv) if packed else raw[lv:])
            if len(page) != h["uncompressed_size"]:
                raise Corrupt("page does not decompress to its declared size")
            h = dict(h, compressed_size=h["uncompressed_size"], is_compressed=False)
            buf, body, end = page, 0, len(page)
        elif h["compressed_size"] != h["uncompressed_size"]:
            raise Unsupported("compressed pages")
        if h["type"] == PAGE_DICTIONARY:
            if dictionary is not None or done:
                raise Corrupt("dictionary page must be the first page of the chunk")
            if h["encoding"] not in (ENC_PLAIN, ENC_PLAIN_DICTIONARY):
                raise Unsupported("dictionary page encoding %d" % h["encoding"])
            dictionary = plain_values(buf, body, end, ptype, h["num_values"], type_length)
            pages.append(("dictionary", h["num_values"]))
            continue
        if h["type"] == PAGE_INDEX:
            continue
        n = h["num_values"]
        if n < 0 or done + n > num_values:
            raise Corrupt("pages hold more values than the chunk metadata says")
        p = body
        if h["type"] == PAGE_DATA:
            if max_def:
                if h["def_encoding"] != ENC_RLE:
                    raise Unsupported("definition level encoding %d" % h["def_encoding"])
                (ln,) = struct.unpack_from("<i", buf, p)
                levels = hybrid_decode(buf, p + 4, p + 4 + ln, 1, n)
                p += 4 + ln
            else:
                levels = np.ones(n, np.uint32)
        elif h["type"] == PAGE_DATA_V2:
            if h.get("rep_bytes", 0):
                raise Unsupported("repetition levels")
            if h.get("is_compressed", True) and h["compressed_size"] != h["uncompressed_size"]:
                raise Unsupported("compressed pages")
            db = h.get("def_bytes", 0)
            levels = hybrid_decode(buf, p, p + db, 1, n) if max_def and db else np.ones(n, np.uint32)
            p += db
        else:
            raise Unsupported("page type %d" % h["type"])
        present = levels == (1 if max_def else 1)
        if max_def == 0:
            present[:] = True
        nn = int(present.sum())
        enc = h["encoding"]
        where = done + np.flatnonzero(present)
        if enc in (ENC_PLAIN_DICTIONARY, ENC_RLE_DICTIONARY):
            if dictionary is None:
                raise Corrupt("Dictionary is missing for Page")
            width = buf[p]
            if width > 32:
                raise Corrupt("dictionary id width %d" % width)
            ids = hybrid_decode(buf, p + 1, end, width, nn) if nn else np.zeros(0, np.uint32)
            if nn and int(ids.max()) >= len(dictionary):
                raise Corrupt("dictionary id out of range")
            ids_all[where] = ids.astype(np.int32)
            if ptype == BYTE_ARRAY:
                for w, i in zip(where, ids):
                    values[w] = dictionary[i]
            else:
                values[where] = np.asarray(dictionary)[ids]
        elif enc == ENC_PLAIN:
            all_dict = False
            v = plain_values(buf, p, end, ptype, nn, type_length)
            if ptype == BYTE_ARRAY:
                for w, x in zip(where, v):
                    values[w] = x
            else:
                values[where] = v
        elif enc == ENC_RLE and ptype == BOOLEAN:
            # ParquetEncoding.java:106-116 with getMaxLevel(VALUES) = 1 for BOOLEAN: 4-byte length + hybrid stream of width 1
            all_dict = False
            (ln,) = struct.unpack_from("<i", buf, p)
            if ln < 0 or p + 4 + ln > end:
                raise Corrupt("RLE boolean stream exceeds the page")
            values[where] = hybrid_decode(buf, p + 4, p + 4 + ln, 1, nn).astype(np.uint8)
        else:
            raise Unsupported("data page encoding %d" % enc)
        nulls[done:done + n] = ~present
        pages.append(("data_v2" if h["type"] == PAGE_DATA_V2 else "data", n, enc, nn))
        done += n
    return values, nulls, {"pages": pages, "all_dictionary": all_dict and dictionary is not None, "dictionary": dictionary, "ids": ids_all}
