// parquet_host.h — host side of ingesting Parquet column chunks: reads the page HEADERS and the RUN HEADERS of the hybrid
// streams of a column chunk (a few bytes per page / per run) and describes where every piece lies. Values are never decoded
// on the CPU: the chunk goes to the device in one DMA and the kernels of parquet.cu expand it there.
//
// Stands in for (paths under /root/reference/lib/trino-parquet/src/main/java/io/trino/parquet/):
//   reader/ParquetColumnChunk.java:54-135      readPageHeader / readAllPages / readDictionaryPage / readDataPageV1 / V2
//   reader/PrimitiveColumnReader.java:268-317  readPageV1 / readPageV2 / initDataReader (levels, then values)
//   ParquetEncoding.java:58-234                PLAIN, RLE, PLAIN_DICTIONARY, RLE_DICTIONARY value readers
// The formats themselves (Thrift compact PageHeader, RLE / bit-packed hybrid) are parquet-format's; the reference takes them
// from the org.apache.parquet classes shaded into io.trino.hive:hive-apache 3.1.2-12 (not in the reference tree).
//
// Subset: flat columns (max definition level 0 or 1, no repetition), UNCOMPRESSED pages, data pages v1 and v2, PLAIN /
// PLAIN_DICTIONARY / RLE_DICTIONARY values (+ RLE booleans), RLE levels; BOOLEAN, INT32, INT64, DOUBLE, FIXED_LEN_BYTE_ARRAY(<= 16)
// and dictionary-encoded BYTE_ARRAY. Everything else is rejected ("not supported ...").
#pragma once
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "zstd_decode.h"

namespace pgpu {
namespace pq {

enum PhysicalType { BOOLEAN = 0, INT32 = 1, INT64 = 2, INT96 = 3, FLOAT = 4, DOUBLE = 5, BYTE_ARRAY = 6, FIXED_LEN_BYTE_ARRAY = 7 };  // parquet.thrift Type
enum Encoding { PLAIN = 0, PLAIN_DICTIONARY = 2, RLE = 3, BIT_PACKED = 4, RLE_DICTIONARY = 8 };                                       // parquet.thrift Encoding
enum PageType { DATA_PAGE = 0, INDEX_PAGE = 1, DICTIONARY_PAGE = 2, DATA_PAGE_V2 = 3 };                                               // parquet.thrift PageType

// One run of a hybrid stream, in the form the device kernels read (24 bytes). `start` counts in the stream's own numbering,
// chunk-wide: rows for definition levels, non-null values for ids / booleans.
struct Run {
    int64_t start;
    uint64_t src;      // bit-packed: BIT offset of the run's first value from the start of the chunk; RLE: the repeated value
    uint32_t count;
    uint32_t widthPacked;  // bits 0..7: bit width; bit 31: bit-packed
};
constexpr uint32_t kPackedFlag = 0x80000000u;

struct DataPage {
    int64_t firstRow = 0, rows = 0;
    int64_t firstValue = 0, nonNull = 0;  // position among the chunk's non-null values
    int encoding = PLAIN;
    bool dictionaryEncoded = false;
    size_t valuesOff = 0, valuesEnd = 0;  // PLAIN fixed width: the values; otherwise the value stream (described by valueRuns)
    size_t runLo = 0, runHi = 0;          // this page's slice of Chunk::valueRuns
};

struct Chunk {
    int64_t rows = 0, nonNull = 0;
    bool hasDictionary = false;
    int64_t dictionaryCount = 0;
    size_t dictionaryOff = 0, dictionaryEnd = 0;  // PLAIN values of the dictionary page
    bool allDictionary = false;                   // every data page is dictionary encoded -> DictionaryBlock
    bool anyDictionary = false;
    std::vector<DataPage> pages;
    std::vector<Run> levelRuns;   // definition levels of the whole chunk (empty when there is no NULL: rows == nonNull)
    std::vector<Run> valueRuns;   // dictionary ids, BOOLEAN values
    bool hasNulls() const { return nonNull != rows; }
};

namespace detail {

struct Reader {
    const uint8_t* d;
    size_t len, pos;
    bool ok = true;
    uint8_t byte() {
        if (pos >= len) { ok = false; return 0; }
        return d[pos++];
    }
    uint64_t varint() {
        uint64_t v = 0;
        for (int s = 0; s < 70; s += 7) {
            const uint8_t c = byte();
            if (!ok) return 0;
            if (s < 64) v |= (uint64_t)(c & 0x7F) << s;
            if (!(c & 0x80)) return v;
        }
        ok = false;
        return 0;
    }
    int64_t zigzag() {
        const uint64_t v = varint();
        return (int64_t)(v >> 1) ^ -(int64_t)(v & 1);
    }
    void advance(uint64_t n) {
        if (n > len - pos) { ok = false; pos = len; return; }
        pos += (size_t)n;
    }
    // Thrift compact protocol: skip one value of wire type t
    void skip(int t, int depth = 0) {
        if (!ok || depth > 16) { ok = false; return; }
        switch (t) {
            case 1: case 2: return;  // booleans live in the field header
            case 3: byte(); return;
            case 4: case 5: case 6: varint(); return;
            case 7: advance(8); return;
            case 8: advance(varint()); return;
            case 9: case 10: {
                const uint8_t h = byte();
                uint64_t n = h >> 4;
                if (n == 15) n = varint();
                for (uint64_t i = 0; i < n && ok; i++) skipElement(h & 15, depth);
                return;
            }
            case 11: {
                const uint64_t n = varint();
                if (!n) return;
                const uint8_t kv = byte();
                for (uint64_t i = 0; i < n && ok; i++) {
                    skipElement(kv >> 4, depth);
                    skipElement(kv & 15, depth);
                }
                return;
            }
            case 12: {
                int16_t last = 0;
                int fid, ft;
                while (field(last, fid, ft)) skip(ft, depth + 1);
                return;
            }
            default: ok = false;
        }
    }
    void skipElement(int t, int depth) {
        if (t == 1 || t == 2) byte();  // booleans inside collections take a byte
        else skip(t, depth + 1);
    }
    // next field header of a struct; false at the stop byte
    bool field(int16_t& last, int& fid, int& type) {
        const uint8_t h = byte();
        if (!ok || h == 0) return false;
        type = h & 15;
        const int delta = h >> 4;
        fid = delta ? last + delta : (int)zigzag();
        last = (int16_t)fid;
        return ok;
    }
};

struct PageHeader {
    int type = -1;
    int64_t uncompressedSize = -1, compressedSize = -1;
    int64_t numValues = -1, numNulls = -1;
    int encoding = -1, defEncoding = RLE, repEncoding = RLE;
    int64_t defBytes = 0, repBytes = 0;
    bool isCompressed = true;  // DataPageHeaderV2 default
};

// parquet.thrift PageHeader {1 type, 2 uncompressed_page_size, 3 compressed_page_size, 4 crc, 5 data_page_header,
// 6 index_page_header, 7 dictionary_page_header, 8 data_page_header_v2}
inline bool readPageHeader(Reader& r, PageHeader& h) {
    int16_t last = 0;
    int fid, t;
    while (r.field(last, fid, t)) {
        if (fid == 1 && t == 5) h.type = (int)r.zigzag();
        else if (fid == 2 && t == 5) h.uncompressedSize = r.zigzag();
        else if (fid == 3 && t == 5) h.compressedSize = r.zigzag();
        else if ((fid == 5 || fid == 7 || fid == 8) && t == 12) {
            int16_t l2 = 0;
            int f2, t2;
            while (r.field(l2, f2, t2)) {
                const bool integer = t2 == 4 || t2 == 5 || t2 == 6;
                if (f2 == 1 && integer) h.numValues = r.zigzag();
                else if (fid != 8 && f2 == 2 && integer) h.encoding = (int)r.zigzag();
                else if (fid == 5 && f2 == 3 && integer) h.defEncoding = (int)r.zigzag();
                else if (fid == 5 && f2 == 4 && integer) h.repEncoding = (int)r.zigzag();
                else if (fid == 8 && f2 == 2 && integer) h.numNulls = r.zigzag();
                else if (fid == 8 && f2 == 4 && integer) h.encoding = (int)r.zigzag();
                else if (fid == 8 && f2 == 5 && integer) h.defBytes = r.zigzag();
                else if (fid == 8 && f2 == 6 && integer) h.repBytes = r.zigzag();
                else if (fid == 8 && f2 == 7 && (t2 == 1 || t2 == 2)) h.isCompressed = t2 == 1;
                else r.skip(t2);
            }
        } else {
            r.skip(t);
        }
    }
    return r.ok && h.type >= 0 && h.uncompressedSize >= 0 && h.compressedSize >= 0;
}

inline int64_t popcountBits(const uint8_t* d, uint64_t bits) {
    int64_t c = 0;
    uint64_t i = 0;
    for (; i + 64 <= bits; i += 64) {
        uint64_t w;
        memcpy(&w, d + i / 8, 8);
        c += __builtin_popcountll(w);
    }
    for (; i < bits; i++) c += (d[i / 8] >> (i & 7)) & 1;
    return c;
}

// Reads the run headers of one hybrid stream d[pos, end) holding `count` values of `width` bits and appends them to `runs`
// (numbered from `start`); *ones = how many values equal 1 (definition levels of width 1). Adjacent RLE runs of the same value
// are merged when `merge` (levels of a column without NULLs collapse into one run per chunk).
inline bool scanHybrid(const uint8_t* d, size_t pos, size_t end, int width, int64_t count, int64_t start, std::vector<Run>& runs, int64_t* ones,
                       bool merge, std::string& why) {
    Reader r{d, end, pos};
    const int valueBytes = (width + 7) / 8;
    int64_t n = 0;
    while (n < count) {
        if (r.pos >= end) { why = "corrupt Parquet page: hybrid stream ends after " + std::to_string(n) + " of " + std::to_string(count) + " values"; return false; }
        const uint64_t h = r.varint();
        if (!r.ok) { why = "corrupt Parquet page: truncated run header"; return false; }
        if (h & 1) {
            const uint64_t groups = h >> 1;
            if (groups == 0 || groups > (1ull << 40)) { why = "corrupt Parquet page: bit-packed run of " + std::to_string(groups) + " groups"; return false; }
            const uint64_t bytes = groups * (uint64_t)width;
            const int64_t take = (int64_t)std::min<uint64_t>(groups * 8, (uint64_t)(count - n));
            // the last group of a stream may be cut short by its writer; what we take must be there
            if (((uint64_t)take * width + 7) / 8 > end - r.pos) { why = "corrupt Parquet page: bit-packed run exceeds the stream"; return false; }
            if (ones) *ones += width == 1 ? popcountBits(d + r.pos, (uint64_t)take) : 0;
            int64_t done = 0;
            while (done < take) {  // Run::count is 32 bits
                const int64_t c = std::min<int64_t>(take - done, 0x40000000);
                runs.push_back(Run{start + n + done, (uint64_t)r.pos * 8 + (uint64_t)done * width, (uint32_t)c, (uint32_t)width | kPackedFlag});
                done += c;
            }
            r.pos += (size_t)std::min<uint64_t>(bytes, end - r.pos);
            n += take;
        } else {
            const uint64_t len = h >> 1;
            if (len == 0) { why = "corrupt Parquet page: empty RLE run"; return false; }
            if ((size_t)valueBytes > end - r.pos) { why = "corrupt Parquet page: RLE run exceeds the stream"; return false; }
            uint64_t v = 0;
            memcpy(&v, d + r.pos, (size_t)valueBytes);
            r.pos += (size_t)valueBytes;
            const int64_t take = (int64_t)std::min<uint64_t>(len, (uint64_t)(count - n));
            if (ones && v == 1) *ones += take;
            int64_t done = 0;
            while (done < take) {
                const int64_t c = std::min<int64_t>(take - done, 0x40000000);
                Run* prev = runs.empty() ? nullptr : &runs.back();
                if (merge && prev && !(prev->widthPacked & kPackedFlag) && prev->src == v && prev->start + prev->count == start + n + done && (int64_t)prev->count + c <= 0x40000000)
                    prev->count += (uint32_t)c;
                else
                    runs.push_back(Run{start + n + done, v, (uint32_t)c, (uint32_t)width});
                done += c;
            }
            n += take;
        }
    }
    return true;
}

inline int32_t le32(const uint8_t* p) {
    int32_t v;
    memcpy(&v, p, 4);
    return v;
}

}  // namespace detail

// bytes per PLAIN value in the file (0 = var-width)
inline int valueWidth(int ptype, int typeLength = 0) {
    return ptype == INT32 ? 4 : ptype == INT64 || ptype == DOUBLE ? 8 : ptype == BOOLEAN ? 1 : ptype == FIXED_LEN_BYTE_ARRAY ? typeLength : 0;
}

namespace detail {

struct PageDesc {
    PageHeader h;
    size_t body = 0, end = 0;
    int64_t firstRow = 0;
};

// What one data page contributes, computed independently of every other page (so pages can be walked in parallel): definition
// level runs numbered by ROW (absolute), value runs numbered from 0 within the page (shifted by the page's first value when the
// chunk is assembled), the non-null count.
struct PageResult {
    bool ok = true;
    std::string why;
    std::vector<Run> levels, values;
    bool scanned = false;  // the page has a definition level stream
    bool plain = false;    // not dictionary encoded
    DataPage pg;
};

inline void appendLevelRun(std::vector<Run>& levels, int64_t at, uint64_t value, int64_t count) {
    while (count > 0) {
        const int64_t c = std::min<int64_t>(count, 0x40000000);
        Run* prev = levels.empty() ? nullptr : &levels.back();
        if (prev && !(prev->widthPacked & kPackedFlag) && prev->src == value && prev->start + prev->count == at && (int64_t)prev->count + c <= 0x40000000)
            prev->count += (uint32_t)c;
        else
            levels.push_back(Run{at, value, (uint32_t)c, 1});
        at += c;
        count -= c;
    }
}

inline void parsePage(const uint8_t* d, const PageDesc& pd, int ptype, int maxDef, int typeLength, bool hasDictionary, PageResult& r) {
    const PageHeader& h = pd.h;
    const size_t end = pd.end;
    auto failed = [&](const std::string& m) { r.ok = false; r.why = m; };
    DataPage& pg = r.pg;
    pg.firstRow = pd.firstRow;
    pg.rows = h.numValues;
    pg.encoding = h.encoding;
    size_t p = pd.body;
    int64_t ones = 0;
    if (h.type == DATA_PAGE) {
        // v1: [repetition levels (absent at max level 0)][4-byte length + hybrid definition levels][values]
        if (maxDef) {
            if (h.defEncoding != RLE) return failed("not supported on the GPU path: definition level encoding " + std::to_string(h.defEncoding));
            if (end - p < 4) return failed("corrupt Parquet page: no room for the level length");
            const int32_t ln = le32(d + p);
            if (ln < 0 || (uint64_t)ln > end - p - 4) return failed("corrupt Parquet page: definition levels exceed the page");
            if (!scanHybrid(d, p + 4, p + 4 + (size_t)ln, 1, pg.rows, pd.firstRow, r.levels, &ones, true, r.why)) { r.ok = false; return; }
            r.scanned = true;
            p += 4 + (size_t)ln;
        } else {
            ones = pg.rows;
        }
    } else {
        if (h.repBytes) return failed("not supported on the GPU path: repetition levels");
        if (h.defBytes < 0 || (uint64_t)h.defBytes > end - p) return failed("corrupt Parquet page: definition levels exceed the page");
        if (maxDef && h.defBytes) {
            if (!scanHybrid(d, p, p + (size_t)h.defBytes, 1, pg.rows, pd.firstRow, r.levels, &ones, true, r.why)) { r.ok = false; return; }
            r.scanned = true;
        } else {
            ones = pg.rows;
        }
        p += (size_t)h.defBytes;
    }
    pg.nonNull = ones;
    if (h.type == DATA_PAGE_V2 && h.numNulls >= 0 && h.numNulls != pg.rows - ones) return failed("corrupt Parquet page: num_nulls disagrees with the definition levels");
    pg.valuesOff = p;
    pg.valuesEnd = end;
    if (h.encoding == PLAIN_DICTIONARY || h.encoding == RLE_DICTIONARY) {
        if (!hasDictionary) return failed("corrupt Parquet chunk: Dictionary is missing for Page");
        pg.dictionaryEncoded = true;
        if (pg.nonNull) {
            if (p >= end) return failed("corrupt Parquet page: no dictionary id width");
            const int width = d[p];
            if (width > 32) return failed("corrupt Parquet page: dictionary id width " + std::to_string(width));
            if (!scanHybrid(d, p + 1, end, width, pg.nonNull, 0, r.values, nullptr, false, r.why)) { r.ok = false; return; }
        }
    } else if (h.encoding == PLAIN) {
        r.plain = true;
        if (ptype == BYTE_ARRAY) return failed("not supported on the GPU path: PLAIN-encoded BYTE_ARRAY data pages (only dictionary-encoded var-width chunks)");
        if (ptype == BOOLEAN) {  // bits, least significant first = one bit-packed run of width 1
            if (((uint64_t)pg.nonNull + 7) / 8 > end - p) return failed("corrupt Parquet page: PLAIN booleans exceed the page");
            int64_t doneV = 0;
            while (doneV < pg.nonNull) {
                const int64_t c = std::min<int64_t>(pg.nonNull - doneV, 0x40000000);
                r.values.push_back(Run{doneV, (uint64_t)p * 8 + (uint64_t)doneV, (uint32_t)c, 1u | kPackedFlag});
                doneV += c;
            }
        } else if ((uint64_t)pg.nonNull * valueWidth(ptype, typeLength) > end - p) {
            return failed("corrupt Parquet page: PLAIN values exceed the page");
        }
    } else if (h.encoding == RLE && ptype == BOOLEAN) {  // ParquetEncoding.java:106-116, max "level" of BOOLEAN values = 1
        r.plain = true;
        if (end - p < 4) return failed("corrupt Parquet page: no room for the RLE length");
        const int32_t ln = le32(d + p);
        if (ln < 0 || (uint64_t)ln > end - p - 4) return failed("corrupt Parquet page: RLE booleans exceed the page");
        if (pg.nonNull && !scanHybrid(d, p + 4, p + 4 + (size_t)ln, 1, pg.nonNull, 0, r.values, nullptr, false, r.why)) { r.ok = false; return; }
    } else {
        return failed("not supported on the GPU path: Parquet data page encoding " + std::to_string(h.encoding));
    }
}

}  // namespace detail

// ------------------------------------------------------------------------------------------------ page decompression
// The reference undoes the codec on the CPU, page by page (lib/trino-parquet/.../reader/PageReader.java:63-99 ->
// ParquetCompressionUtils.decompress :43-70, airlift aircompressor). So does this ingest: byte-serial LZ77 decoding stays on the
// host (one thread per column chunk), everything after it — levels, dictionary ids, NULL expansion — on the device.
enum Codec { UNCOMPRESSED = 0, SNAPPY = 1, GZIP = 2, LZO = 3, BROTLI = 4, LZ4 = 5, ZSTD = 6, LZ4_RAW = 7 };  // parquet.thrift CompressionCodec

// Raw Snappy block (format_description.txt of the Snappy sources): varint uncompressed length, then elements tagged by the low
// two bits of their first byte: 00 literal, 01 / 10 / 11 copy with a 1- / 2- / 4-byte offset. Copies may overlap their output.
inline bool snappyDecompress(const uint8_t* in, size_t n, uint8_t* out, size_t outLen) {
    size_t ip = 0, op = 0;
    uint64_t declared = 0;
    for (int shift = 0;; shift += 7) {
        if (ip >= n || shift > 35) return false;
        const uint8_t b = in[ip++];
        declared |= (uint64_t)(b & 0x7f) << shift;
        if (!(b & 0x80)) break;
    }
    if (declared != outLen) return false;
    while (ip < n) {
        const uint8_t tag = in[ip++];
        size_t len, offset;
        if ((tag & 3) == 0) {
            len = (size_t)(tag >> 2) + 1;
            if (len > 60) {
                const int extra = (int)len - 60;  // 1..4 length bytes, little endian
                if (ip + (size_t)extra > n) return false;
                len = 0;
                for (int k = 0; k < extra; k++) len |= (size_t)in[ip + k] << (8 * k);
                len += 1;
                ip += (size_t)extra;
            }
            if (len > n - ip || len > outLen - op) return false;
            memcpy(out + op, in + ip, len);
            ip += len;
            op += len;
            continue;
        }
        if ((tag & 3) == 1) {
            if (ip >= n) return false;
            len = (size_t)((tag >> 2) & 7) + 4;
            offset = ((size_t)(tag >> 5) << 8) | in[ip++];
        } else if ((tag & 3) == 2) {
            if (ip + 2 > n) return false;
            len = (size_t)(tag >> 2) + 1;
            offset = (size_t)in[ip] | ((size_t)in[ip + 1] << 8);
            ip += 2;
        } else {
            if (ip + 4 > n) return false;
            len = (size_t)(tag >> 2) + 1;
            offset = (size_t)in[ip] | ((size_t)in[ip + 1] << 8) | ((size_t)in[ip + 2] << 16) | ((size_t)in[ip + 3] << 24);
            ip += 4;
        }
        if (offset == 0 || offset > op || len > outLen - op) return false;
        if (offset >= len) memcpy(out + op, out + op - offset, len);
        else for (size_t k = 0; k < len; k++) out[op + k] = out[op - offset + k];  // overlapping: byte by byte
        op += len;
    }
    return op == outLen;
}

// LZ4 block format (lz4_Block_format.md): sequences of [token | literal length bytes | literals | offset LE16 | match length bytes];
// the last sequence ends after its literals.
// `produced` given: the block may fill less than outLen (a chunk of a Hadoop frame, whose own size is not stored); else it must fill it
inline bool lz4Decompress(const uint8_t* in, size_t n, uint8_t* out, size_t outLen, size_t* produced = nullptr) {
    size_t ip = 0, op = 0;
    while (ip < n) {
        const uint8_t token = in[ip++];
        size_t lit = token >> 4;
        if (lit == 15) {
            uint8_t b;
            do {
                if (ip >= n) return false;
                b = in[ip++];
                lit += b;
            } while (b == 255);
        }
        if (lit > n - ip || lit > outLen - op) return false;
        memcpy(out + op, in + ip, lit);
        ip += lit;
        op += lit;
        if (ip >= n) break;  // last sequence: literals only
        if (ip + 2 > n) return false;
        const size_t offset = (size_t)in[ip] | ((size_t)in[ip + 1] << 8);
        ip += 2;
        size_t len = (size_t)(token & 15) + 4;
        if ((token & 15) == 15) {
            uint8_t b;
            do {
                if (ip >= n) return false;
                b = in[ip++];
                len += b;
            } while (b == 255);
        }
        if (offset == 0 || offset > op || len > outLen - op) return false;
        if (offset >= len) memcpy(out + op, out + op - offset, len);
        else for (size_t k = 0; k < len; k++) out[op + k] = out[op - offset + k];
        op += len;
    }
    if (produced) {
        *produced = op;
        return true;
    }
    return op == outLen;
}

// The deprecated Parquet codec LZ4 (5) as Hadoop's Lz4Codec frames it and the reference reads it (ParquetCompressionUtils.java:
// 104-135 decompressFramed): [uncompressed length of a block, big endian][compressed length of a chunk, big endian][LZ4 block]...,
// a block may come in several chunks.
inline bool lz4HadoopDecompress(const uint8_t* in, size_t n, uint8_t* out, size_t outLen) {
    auto be32 = [&](size_t at) { return ((size_t)in[at] << 24) | ((size_t)in[at + 1] << 16) | ((size_t)in[at + 2] << 8) | (size_t)in[at + 3]; };
    size_t ip = 0, op = 0, blockEnd = 0;
    while (op < outLen) {
        if (op == blockEnd) {
            if (ip + 4 > n) return false;
            blockEnd += be32(ip);
            ip += 4;
            if (blockEnd > outLen) return false;
        }
        if (ip + 4 > n) return false;
        const size_t chunk = be32(ip);
        ip += 4;
        if (chunk > n - ip) return false;
        size_t got = 0;
        if (!lz4Decompress(in + ip, chunk, out + op, blockEnd - op, &got) || got == 0) return false;
        op += got;
        ip += chunk;
    }
    return op == outLen && ip == n;  // nothing may follow the last chunk
}

// DEFLATE (RFC 1951) inside a gzip member (RFC 1952): what Hadoop's GzipCodec — the reference's GZIP pages — writes.
// Canonical Huffman codes are decoded bit by bit from per-length counts and the symbols sorted by (length, value).
struct Inflater {
    const uint8_t* in;
    size_t n, ip = 0;
    uint64_t hold = 0;
    int nbits = 0;
    bool ok = true;
    void refill() {  // at least 56 bits in `hold` while input lasts
        while (nbits <= 56 && ip < n) {
            hold |= (uint64_t)in[ip++] << nbits;
            nbits += 8;
        }
    }
    uint32_t bits(int need) {
        if (nbits < need) {
            refill();
            if (nbits < need) { ok = false; return 0; }
        }
        const uint32_t v = (uint32_t)(hold & ((1ull << need) - 1ull));
        hold >>= need;
        nbits -= need;
        return v;
    }
    static constexpr int kFast = 10;  // codes of up to 10 bits are decoded with one table look-up
    struct Code {
        uint16_t count[16];
        uint16_t symbol[288];
        uint16_t fast[1 << kFast];  // indexed by the next kFast stream bits: symbol << 4 | code length, 0 = longer code
    };
    // false when the lengths over-subscribe the code space
    static bool build(Code& c, const uint8_t* lengths, int nsym) {
        memset(c.count, 0, sizeof c.count);
        for (int k = 0; k < nsym; k++) c.count[lengths[k]]++;
        int left = 1;
        for (int len = 1; len <= 15; len++) {
            left = (left << 1) - c.count[len];
            if (left < 0) return false;
        }
        uint16_t first[16], next[16];
        first[1] = 0;
        for (int len = 1; len < 15; len++) first[len + 1] = (uint16_t)(first[len] + c.count[len]);
        for (int k = 0; k < nsym; k++)
            if (lengths[k]) c.symbol[first[lengths[k]]++] = (uint16_t)k;
        // canonical codes (RFC 1951 3.2.2), bit-reversed because the stream delivers a code's most significant bit first
        memset(c.fast, 0, sizeof c.fast);
        int code = 0;
        for (int len = 1; len <= 15; len++) {  // first code of every length; unused symbols (length 0) have none
            next[len] = (uint16_t)code;
            code = (code + c.count[len]) << 1;
        }
        for (int k = 0; k < nsym; k++) {
            const int len = lengths[k];
            if (!len) continue;
            const int cw = next[len]++;
            if (len > kFast) continue;
            int rev = 0;
            for (int b = 0; b < len; b++) rev |= ((cw >> b) & 1) << (len - 1 - b);
            for (int fill = rev; fill < (1 << kFast); fill += 1 << len) c.fast[fill] = (uint16_t)((k << 4) | len);
        }
        return true;
    }
    int decodeSlow(const Code& c) {
        int code = 0, first = 0, index = 0;
        for (int len = 1; len <= 15; len++) {
            code |= (int)bits(1);
            if (!ok) return -1;
            const int count = c.count[len];
            if (code - count < first) return c.symbol[index + (code - first)];
            index += count;
            first = (first + count) << 1;
            code <<= 1;
        }
        ok = false;
        return -1;
    }
    int decode(const Code& c) {
        if (nbits < 15) refill();
        const uint16_t e = c.fast[hold & ((1u << kFast) - 1u)];
        const int len = e & 15;
        if (len && len <= nbits) {
            hold >>= len;
            nbits -= len;
            return e >> 4;
        }
        return decodeSlow(c);
    }
    bool codes(const Code& lit, const Code& dist, uint8_t* out, size_t outLen, size_t& op) {
        static const uint16_t lbase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
        static const uint8_t lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
        static const uint16_t dbase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
        static const uint8_t dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
        for (;;) {
            const int sym = decode(lit);
            if (sym < 0) return false;
            if (sym < 256) {
                if (op >= outLen) return false;
                out[op++] = (uint8_t)sym;
            } else if (sym == 256) {
                return true;
            } else {
                if (sym > 285) return false;
                const size_t len = lbase[sym - 257] + bits(lext[sym - 257]);
                const int ds = decode(dist);
                if (ds < 0 || ds > 29) return false;
                const size_t d = dbase[ds] + bits(dext[ds]);
                if (!ok || d > op || len > outLen - op) return false;
                if (d >= len) memcpy(out + op, out + op - d, len);
                else for (size_t k = 0; k < len; k++) out[op + k] = out[op - d + k];
                op += len;
            }
        }
    }
    bool inflate(uint8_t* out, size_t outLen, size_t& op) {
        for (;;) {
            const bool last = bits(1) != 0;
            const uint32_t type = bits(2);
            if (!ok) return false;
            if (type == 0) {  // stored: the rest of the current byte is skipped, whole bytes that were read ahead go back
                ip -= (size_t)(nbits / 8);
                hold = 0;
                nbits = 0;
                if (ip + 4 > n) return false;
                const size_t len = (size_t)in[ip] | ((size_t)in[ip + 1] << 8), nlen = (size_t)in[ip + 2] | ((size_t)in[ip + 3] << 8);
                ip += 4;
                if ((len ^ nlen) != 0xffff || len > n - ip || len > outLen - op) return false;
                memcpy(out + op, in + ip, len);
                ip += len;
                op += len;
            } else if (type == 1) {  // fixed codes
                uint8_t lengths[288 + 30];
                for (int k = 0; k < 288; k++) lengths[k] = k < 144 ? 8 : k < 256 ? 9 : k < 280 ? 7 : 8;
                for (int k = 0; k < 30; k++) lengths[288 + k] = 5;
                Code lit, dist;
                build(lit, lengths, 288);
                build(dist, lengths + 288, 30);
                if (!codes(lit, dist, out, outLen, op)) return false;
            } else if (type == 2) {  // dynamic codes
                static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
                const int nlen = (int)bits(5) + 257, ndist = (int)bits(5) + 1, ncode = (int)bits(4) + 4;
                if (!ok || nlen > 286 || ndist > 30) return false;
                uint8_t lengths[288 + 32];
                memset(lengths, 0, sizeof lengths);
                for (int k = 0; k < ncode; k++) lengths[order[k]] = (uint8_t)bits(3);
                Code lencode;
                if (!ok || !build(lencode, lengths, 19)) return false;
                uint8_t ll[288 + 32];
                int idx = 0;
                while (idx < nlen + ndist) {
                    const int sym = decode(lencode);
                    if (sym < 0) return false;
                    if (sym < 16) {
                        ll[idx++] = (uint8_t)sym;
                        continue;
                    }
                    uint8_t prev = 0;
                    int rep;
                    if (sym == 16) {
                        if (idx == 0) return false;
                        prev = ll[idx - 1];
                        rep = 3 + (int)bits(2);
                    } else if (sym == 17) {
                        rep = 3 + (int)bits(3);
                    } else {
                        rep = 11 + (int)bits(7);
                    }
                    if (!ok || idx + rep > nlen + ndist) return false;
                    while (rep--) ll[idx++] = prev;
                }
                if (ll[256] == 0) return false;  // no end-of-block code
                Code lit, dist;
                if (!build(lit, ll, nlen) || !build(dist, ll + nlen, ndist)) return false;
                if (!codes(lit, dist, out, outLen, op)) return false;
            } else {
                return false;
            }
            if (last) {
                ip -= (size_t)(nbits / 8);  // bytes read ahead but not used belong to what follows the stream (the gzip trailer)
                nbits = 0;
                hold = 0;
                return ok;
            }
        }
    }
};

inline uint32_t crc32Of(const uint8_t* d, size_t n) {
    static uint32_t table[256];
    static bool ready = [] {
        for (uint32_t k = 0; k < 256; k++) {
            uint32_t c = k;
            for (int j = 0; j < 8; j++) c = (c & 1) ? 0xedb88320u ^ (c >> 1) : c >> 1;
            table[k] = c;
        }
        return true;
    }();
    (void)ready;
    uint32_t c = 0xffffffffu;
    for (size_t k = 0; k < n; k++) c = table[(c ^ d[k]) & 0xff] ^ (c >> 8);
    return c ^ 0xffffffffu;
}

// one or more gzip members -> exactly outLen bytes; CRC-32 and ISIZE of every member are checked
inline bool gzipDecompress(const uint8_t* in, size_t n, uint8_t* out, size_t outLen) {
    size_t ip = 0, op = 0;
    while (ip < n) {
        if (n - ip < 18 || in[ip] != 0x1f || in[ip + 1] != 0x8b || in[ip + 2] != 8) return false;
        const uint8_t flg = in[ip + 3];
        ip += 10;
        if (flg & 4) {  // FEXTRA
            if (ip + 2 > n) return false;
            const size_t x = (size_t)in[ip] | ((size_t)in[ip + 1] << 8);
            ip += 2 + x;
        }
        for (int f = 8; f <= 16; f <<= 1)  // FNAME, FCOMMENT: zero-terminated
            if (flg & f) {
                while (ip < n && in[ip]) ip++;
                ip++;
            }
        if (flg & 2) ip += 2;  // FHCRC
        if (ip >= n) return false;
        Inflater z{in + ip, n - ip};
        const size_t start = op;
        if (!z.inflate(out, outLen, op)) return false;
        ip += z.ip;  // whole bytes consumed: the trailer is byte aligned
        if (n - ip < 8) return false;
        uint32_t crc, isize;
        memcpy(&crc, in + ip, 4);
        memcpy(&isize, in + ip + 4, 4);
        ip += 8;
        if (isize != (uint32_t)(op - start) || crc != crc32Of(out + start, op - start)) return false;
    }
    return op == outLen;
}

// Thrift compact protocol, just enough to write a page header
struct HeaderWriter {
    std::vector<uint8_t>& out;
    int16_t last = 0;
    void varint(uint64_t v) {
        while (v >= 0x80) { out.push_back((uint8_t)(v | 0x80)); v >>= 7; }
        out.push_back((uint8_t)v);
    }
    void fieldHeader(int fid, int type) {
        const int delta = fid - last;
        if (delta > 0 && delta <= 15) out.push_back((uint8_t)((delta << 4) | type));
        else { out.push_back((uint8_t)type); varint(((uint64_t)(int64_t)fid << 1) ^ (uint64_t)((int64_t)fid >> 63)); }
        last = (int16_t)fid;
    }
    void i32(int fid, int64_t v) {
        fieldHeader(fid, 5);
        varint(((uint64_t)v << 1) ^ (uint64_t)(v >> 63));
    }
    void boolean(int fid, bool v) { fieldHeader(fid, v ? 1 : 2); }
    void stop() { out.push_back(0); }
};

// A column chunk whose pages are compressed -> the same chunk with UNCOMPRESSED pages (fresh headers that carry what parseChunk
// reads: sizes, value counts, encodings, level byte lengths), which the rest of the ingest takes as it is.
inline bool inflateChunk(const uint8_t* d, size_t len, int codec, int64_t numValues, std::vector<uint8_t>& out, std::string& why, int threads = 1) {
    using namespace detail;
    if (codec != SNAPPY && codec != LZ4_RAW && codec != GZIP && codec != LZ4 && codec != ZSTD) {
        why = "not supported on the GPU path: Parquet compression codec " + std::to_string(codec) + " (UNCOMPRESSED, SNAPPY, GZIP, LZ4, ZSTD and LZ4_RAW pages are read)";
        return false;
    }
    out.clear();
    out.reserve(len * 3);
    struct Task {
        size_t src, srcLen, dst, dstLen;  // the compressed part of one page -> its place in `out`
        bool packed;
    };
    std::vector<Task> tasks;  // pages are independent: headers are written in order here, bodies are inflated in parallel below
    size_t pos = 0;
    int64_t done = 0;
    while (done < numValues) {
        if (pos >= len) { why = "corrupt Parquet chunk: ends after " + std::to_string(done) + " of " + std::to_string(numValues) + " values"; return false; }
        Reader r{d, len, pos};
        PageHeader h;
        if (!readPageHeader(r, h)) { why = "corrupt Parquet chunk: unreadable page header at byte " + std::to_string(pos); return false; }
        const size_t body = r.pos;
        if ((uint64_t)h.compressedSize > len - body) { why = "corrupt Parquet chunk: page exceeds the chunk"; return false; }
        pos = body + (size_t)h.compressedSize;
        if (h.type == INDEX_PAGE) continue;
        if (h.type != DATA_PAGE && h.type != DATA_PAGE_V2 && h.type != DICTIONARY_PAGE) { why = "not supported on the GPU path: Parquet page type " + std::to_string(h.type); return false; }
        if (h.uncompressedSize > (int64_t)1 << 31) { why = "corrupt Parquet chunk: page size"; return false; }
        // data page v2 keeps its levels outside the compressed part (parquet-format DataPageHeaderV2)
        size_t levels = 0;
        bool packed = true;
        if (h.type == DATA_PAGE_V2) {
            if (h.defBytes < 0 || h.repBytes < 0 || h.defBytes + h.repBytes > h.compressedSize || h.defBytes + h.repBytes > h.uncompressedSize) { why = "corrupt Parquet chunk: level byte lengths"; return false; }
            levels = (size_t)(h.defBytes + h.repBytes);
            packed = h.isCompressed;
        }
        HeaderWriter w{out};
        w.i32(1, h.type);
        w.i32(2, h.uncompressedSize);
        w.i32(3, h.uncompressedSize);
        if (h.type == DATA_PAGE) {
            w.fieldHeader(5, 12);
            HeaderWriter s{out};
            s.i32(1, h.numValues); s.i32(2, h.encoding); s.i32(3, h.defEncoding); s.i32(4, h.repEncoding);
            s.stop();
        } else if (h.type == DICTIONARY_PAGE) {
            w.fieldHeader(7, 12);
            HeaderWriter s{out};
            s.i32(1, h.numValues); s.i32(2, h.encoding);
            s.stop();
        } else {
            w.fieldHeader(8, 12);
            HeaderWriter s{out};
            s.i32(1, h.numValues); s.i32(2, h.numNulls < 0 ? 0 : h.numNulls); s.i32(3, h.numValues); s.i32(4, h.encoding); s.i32(5, h.defBytes); s.i32(6, h.repBytes);
            s.boolean(7, false);
            s.stop();
        }
        w.stop();
        const size_t at = out.size();
        out.resize(at + (size_t)h.uncompressedSize);
        memcpy(out.data() + at, d + body, levels);
        tasks.push_back(Task{body + levels, (size_t)h.compressedSize - levels, at + levels, (size_t)h.uncompressedSize - levels, packed});
        if (!packed && tasks.back().srcLen != tasks.back().dstLen) { why = "corrupt Parquet chunk: page sizes"; return false; }
        if (h.type != DICTIONARY_PAGE) {
            if (h.numValues < 0 || h.numValues > numValues - done) { why = "corrupt Parquet chunk: pages hold more values than the metadata says"; return false; }
            done += h.numValues;
        }
    }
    std::vector<char> bad(tasks.size(), 0);
    auto run = [&](size_t lo, size_t hi) {
        for (size_t k = lo; k < hi; k++) {
            const Task& t = tasks[k];
            uint8_t* dst = out.data() + t.dst;
            bool ok = true;
            if (!t.packed) memcpy(dst, d + t.src, t.dstLen);
            else if (t.dstLen)  // (an empty page may be stored as zero bytes)
                ok = codec == SNAPPY ? snappyDecompress(d + t.src, t.srcLen, dst, t.dstLen) : codec == GZIP ? gzipDecompress(d + t.src, t.srcLen, dst, t.dstLen)
                     : codec == LZ4 ? lz4HadoopDecompress(d + t.src, t.srcLen, dst, t.dstLen) : codec == ZSTD ? zstd::decompress(d + t.src, t.srcLen, dst, t.dstLen)
                                    : lz4Decompress(d + t.src, t.srcLen, dst, t.dstLen);
            bad[k] = ok ? 0 : 1;
        }
    };
    const size_t nt = std::min<size_t>((size_t)std::max(threads, 1), std::max<size_t>(tasks.size() / 2, 1));
    if (nt > 1) {
        std::vector<std::thread> workers;
        for (size_t t = 1; t < nt; t++) workers.emplace_back(run, tasks.size() * t / nt, tasks.size() * (t + 1) / nt);
        run(0, tasks.size() / nt);
        for (auto& w : workers) w.join();
    } else {
        run(0, tasks.size());
    }
    for (char b : bad)
        if (b) { why = std::string("corrupt Parquet chunk: ") + (codec == SNAPPY ? "SNAPPY" : codec == GZIP ? "GZIP" : codec == LZ4 ? "LZ4" : codec == ZSTD ? "ZSTD" : "LZ4_RAW") + " page does not decompress to its declared size"; return false; }
    return true;
}

// Describes one column chunk. `why` starts with "not supported" for rejections of well-formed input. Three steps: the page
// headers in file order (each says where the next one lies), the run headers of every data page (pages are independent:
// `threads` > 1 walks them in parallel), then the pages' pieces put together in order (non-null prefix sums fix where each
// page's values land).
inline bool parseChunk(const uint8_t* d, size_t len, int ptype, int maxDef, int64_t numValues, Chunk& out, std::string& why, int typeLength = 0, int threads = 1) {
    using namespace detail;
    if (maxDef < 0 || maxDef > 1) { why = "not supported on the GPU path: nested Parquet columns (max definition level " + std::to_string(maxDef) + ")"; return false; }
    if (!(ptype == BOOLEAN || ptype == INT32 || ptype == INT64 || ptype == DOUBLE || ptype == BYTE_ARRAY || ptype == FIXED_LEN_BYTE_ARRAY)) {
        why = "not supported on the GPU path: Parquet physical type " + std::to_string(ptype);
        return false;
    }
    if (ptype == FIXED_LEN_BYTE_ARRAY && (typeLength < 1 || typeLength > 16)) {
        why = "not supported on the GPU path: FIXED_LEN_BYTE_ARRAY(" + std::to_string(typeLength) + ") (only DECIMALs of up to 16 bytes)";
        return false;
    }
    if (numValues < 0 || (!d && len)) { why = "corrupt Parquet chunk: bad arguments"; return false; }
    out = Chunk();
    out.rows = numValues;
    // 1. page headers
    std::vector<PageDesc> descs;
    size_t pos = 0;
    int64_t done = 0;
    while (done < numValues) {
        if (pos >= len) { why = "corrupt Parquet chunk: ends after " + std::to_string(done) + " of " + std::to_string(numValues) + " values"; return false; }
        Reader r{d, len, pos};
        PageDesc pd;
        PageHeader& h = pd.h;
        if (!readPageHeader(r, h)) { why = "corrupt Parquet chunk: unreadable page header at byte " + std::to_string(pos); return false; }
        if (h.compressedSize != h.uncompressedSize) { why = "not supported on the GPU path: compressed Parquet pages (decompress before handing the chunk over)"; return false; }
        pd.body = r.pos;
        if ((uint64_t)h.compressedSize > len - pd.body) { why = "corrupt Parquet chunk: page exceeds the chunk"; return false; }
        pd.end = pd.body + (size_t)h.compressedSize;
        pos = pd.end;
        if (h.type == INDEX_PAGE) continue;
        if (h.type == DICTIONARY_PAGE) {
            if (out.hasDictionary || done) { why = "corrupt Parquet chunk: dictionary page must be the first page"; return false; }
            if (h.encoding != PLAIN && h.encoding != PLAIN_DICTIONARY) { why = "not supported on the GPU path: dictionary page encoding " + std::to_string(h.encoding); return false; }
            if (ptype == BOOLEAN) { why = "not supported on the GPU path: dictionary-encoded BOOLEAN"; return false; }
            if (h.numValues < 0 || h.numValues > 0x7fffffff) { why = "corrupt Parquet chunk: dictionary size"; return false; }
            const int w = valueWidth(ptype, typeLength);
            if (w && (uint64_t)h.numValues * w > pd.end - pd.body) { why = "corrupt Parquet chunk: dictionary values exceed the page"; return false; }
            out.hasDictionary = true;
            out.dictionaryCount = h.numValues;
            out.dictionaryOff = pd.body;
            out.dictionaryEnd = pd.end;
            continue;
        }
        if (h.type != DATA_PAGE && h.type != DATA_PAGE_V2) { why = "not supported on the GPU path: Parquet page type " + std::to_string(h.type); return false; }
        if (h.numValues < 0 || h.numValues > numValues - done) { why = "corrupt Parquet chunk: pages hold more values than the metadata says"; return false; }
        pd.firstRow = done;
        done += h.numValues;
        descs.push_back(pd);
    }
    // 2. run headers, page by page
    std::vector<PageResult> res(descs.size());
    auto walk = [&](size_t lo, size_t hi) {
        for (size_t k = lo; k < hi; k++) parsePage(d, descs[k], ptype, maxDef, typeLength, out.hasDictionary, res[k]);
    };
    const size_t np = descs.size();
    const size_t nt = std::min<size_t>((size_t)std::max(threads, 1), np / 32);  // at least 32 pages (~ 0.3 ms of walking) per thread: below that a thread costs more than it saves
    if (nt > 1) {
        std::vector<std::thread> workers;
        for (size_t t = 1; t < nt; t++) workers.emplace_back(walk, np * t / nt, np * (t + 1) / nt);
        walk(0, np / nt);
        for (auto& w : workers) w.join();
    } else {
        walk(0, np);
    }
    // 3. assemble in page order
    std::vector<Run> levels;
    int64_t values = 0;
    bool anyPlain = false;
    for (size_t k = 0; k < np; k++) {
        PageResult& r = res[k];
        if (!r.ok) { why = r.why.empty() ? "corrupt Parquet page" : r.why; return false; }
        DataPage& pg = r.pg;
        if (r.scanned) {
            for (const Run& run : r.levels) {
                if (run.widthPacked & kPackedFlag) levels.push_back(run);
                else appendLevelRun(levels, run.start, run.src, run.count);
            }
        } else {
            appendLevelRun(levels, pg.firstRow, 1, pg.rows);  // no level stream: all present; keeps the chunk-wide stream gap-free
        }
        pg.firstValue = values;
        pg.runLo = out.valueRuns.size();
        for (Run run : r.values) {
            run.start += values;
            out.valueRuns.push_back(run);
        }
        pg.runHi = out.valueRuns.size();
        out.anyDictionary |= pg.dictionaryEncoded;
        anyPlain |= r.plain;
        values += pg.nonNull;
        out.pages.push_back(pg);
    }
    out.nonNull = values;
    out.allDictionary = out.anyDictionary && !anyPlain;
    if (ptype == BYTE_ARRAY && !out.allDictionary && out.rows) {
        why = "not supported on the GPU path: var-width Parquet chunk that is not dictionary encoded throughout";
        return false;
    }
    if (out.hasNulls()) out.levelRuns.swap(levels);
    return true;
}

// PLAIN BYTE_ARRAY dictionary page -> offsets[count + 1] + the value bytes (4-byte length prefixes dropped). `maxChars` > 0
// truncates every value to that many UTF-8 code points, as BinaryColumnReader does for a bounded VARCHAR
// (reader/BinaryColumnReader.java:44-58, spi/type/Varchars.truncateToLength).
inline bool binaryDictionary(const uint8_t* d, size_t off, size_t end, int64_t count, int maxChars, std::vector<int32_t>& offsets, std::vector<uint8_t>& bytes, std::string& why) {
    offsets.assign(1, 0);
    bytes.clear();
    size_t p = off;
    for (int64_t i = 0; i < count; i++) {
        if (end - p < 4) { why = "corrupt Parquet dictionary page: truncated length"; return false; }
        const int32_t n = detail::le32(d + p);
        p += 4;
        if (n < 0 || (uint64_t)n > end - p) { why = "corrupt Parquet dictionary page: value exceeds the page"; return false; }
        size_t keep = (size_t)n;
        if (maxChars > 0) {
            int chars = 0;
            size_t k = 0;
            while (k < (size_t)n) {
                if ((d[p + k] & 0xC0) != 0x80) {  // a code point starts here
                    if (chars == maxChars) break;
                    chars++;
                }
                k++;
            }
            keep = k;
        }
        bytes.insert(bytes.end(), d + p, d + p + keep);
        if (bytes.size() > 0x7fffffff) { why = "Parquet dictionary page larger than 2 GB"; return false; }
        offsets.push_back((int32_t)bytes.size());
        p += (size_t)n;
    }
    return true;
}

}  // namespace pq
}  // namespace pgpu
