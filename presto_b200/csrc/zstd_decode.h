// zstd_decode.h — Zstandard frames (RFC 8878) -> bytes, on the host. The reference reads ZSTD Parquet pages with airlift's
// ZstdDecompressor on the CPU (lib/trino-parquet/src/main/java/io/trino/parquet/ParquetCompressionUtils.java:76-81); the
// aircompressor sources are not in /root/reference, so this follows the published format: frame header, raw / RLE / compressed
// blocks, Huffman-coded literals (direct or FSE-compressed weights, 1 or 4 streams, tree reuse), FSE-coded sequences with the
// predefined / RLE / described / repeated tables and the three repeat offsets. No dictionaries (Parquet pages carry none); the
// optional content checksum is skipped, the declared sizes are enforced. Written to be safe on damaged input: every read is
// bounds checked, a failure is `false`, never a fault.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>

namespace pgpu {
namespace zstd {

inline int highBit(uint32_t v) { return 31 - __builtin_clz(v); }  // v != 0

// ---- forward bit reader (FSE table descriptions): least significant bit first
struct Forward {
    const uint8_t* p;
    size_t n;
    size_t bit = 0;
    uint32_t peek(int k) const {  // k <= 24; bits past the end read as zero
        uint64_t w = 0;
        const size_t at = bit >> 3;
        for (int b = 0; b < 5; b++)
            if (at + (size_t)b < n) w |= (uint64_t)p[at + b] << (8 * b);
        return (uint32_t)((w >> (bit & 7)) & ((1u << k) - 1u));
    }
    void skip(int k) { bit += (size_t)k; }
    uint32_t read(int k) {
        const uint32_t v = peek(k);
        bit += (size_t)k;
        return v;
    }
    bool overrun() const { return bit > n * 8; }
    size_t bytesUsed() const { return (bit + 7) >> 3; }
};

// ---- backward bit reader (Huffman streams, sequences): the stream ends with a 1 bit marker in its last byte and is read from
// there towards its first byte. `pos` = unread bits; it goes negative when a reader asks for more than the stream holds.
struct Backward {
    const uint8_t* p = nullptr;
    size_t n = 0;
    int64_t pos = 0;
    bool init(const uint8_t* d, size_t len) {
        p = d;
        n = len;
        if (!len || !d[len - 1]) return false;
        pos = (int64_t)(len - 1) * 8 + highBit(d[len - 1]);
        return true;
    }
    uint64_t bitsAt(int64_t at, int k) const {  // k <= 32 bits starting at bit index `at` (may be negative: zeros)
        uint64_t w = 0;
        const int64_t byte = at >= 0 ? at >> 3 : -((-at + 7) >> 3);
        const int shift = (int)(at - byte * 8);
        if (byte >= 0 && (size_t)byte + 8 <= n) {  // the common case: one unaligned load
            memcpy(&w, p + byte, 8);
            return (w >> shift) & ((1ull << k) - 1ull);
        }
        for (int b = 0; b < 6; b++) {
            const int64_t q = byte + b;
            if (q >= 0 && (size_t)q < n) w |= (uint64_t)p[q] << (8 * b);
        }
        return (w >> shift) & ((1ull << k) - 1ull);
    }
    uint32_t peek(int k) const { return k ? (uint32_t)bitsAt(pos - k, k) : 0u; }
    uint32_t read(int k) {
        pos -= k;
        return k ? (uint32_t)bitsAt(pos, k) : 0u;
    }
};

// ---- FSE
struct FseCell {
    uint16_t newState;
    uint8_t symbol, bits;
};
struct FseTable {
    int log = 0;
    std::vector<FseCell> cell;
};

// table of 1 << log cells from normalized counts (-1 = "less than one": one cell at the end of the table)
inline bool fseBuild(FseTable& t, const int16_t* norm, int nsym, int log) {
    const int size = 1 << log;
    t.log = log;
    t.cell.assign((size_t)size, FseCell{0, 0, 0});
    std::vector<uint16_t> next((size_t)nsym, 0);
    int high = size - 1;
    for (int s = 0; s < nsym; s++)
        if (norm[s] == -1) {
            if (high < 0) return false;
            t.cell[(size_t)high--].symbol = (uint8_t)s;
            next[(size_t)s] = 1;
        }
    const int step = (size >> 1) + (size >> 3) + 3, mask = size - 1;
    int pos = 0;
    for (int s = 0; s < nsym; s++) {
        if (norm[s] <= 0) continue;
        next[(size_t)s] = (uint16_t)norm[s];
        for (int i = 0; i < norm[s]; i++) {
            t.cell[(size_t)pos].symbol = (uint8_t)s;
            do pos = (pos + step) & mask; while (pos > high);
        }
    }
    if (pos != 0) return false;  // the counts did not add up to the table size
    for (int u = 0; u < size; u++) {
        const int s = t.cell[(size_t)u].symbol;
        const uint32_t x = next[(size_t)s]++;
        if (!x) return false;
        const int bits = log - highBit(x);
        t.cell[(size_t)u].bits = (uint8_t)bits;
        t.cell[(size_t)u].newState = (uint16_t)((x << bits) - (uint32_t)size);
    }
    return true;
}

inline void fseRle(FseTable& t, uint8_t symbol) {
    t.log = 0;
    t.cell.assign(1, FseCell{0, symbol, 0});
}

// RFC 8878 4.1.1: the normalized counts of an FSE table; *used = bytes of the description
inline bool fseReadCounts(const uint8_t* d, size_t n, int maxLog, int maxSym, int16_t* norm, int* nsym, int* log, size_t* used) {
    Forward f{d, n};
    const int al = (int)f.read(4) + 5;
    if (al > maxLog) return false;
    int remaining = (1 << al) + 1, threshold = 1 << al, nbBits = al + 1, sym = 0;
    while (remaining > 1 && sym <= maxSym) {
        const int max = (2 * threshold - 1) - remaining;
        const uint32_t bits = f.peek(nbBits);
        int count;
        if ((int)(bits & (uint32_t)(threshold - 1)) < max) {
            count = (int)(bits & (uint32_t)(threshold - 1));
            f.skip(nbBits - 1);
        } else {
            count = (int)(bits & (uint32_t)(2 * threshold - 1));
            if (count >= threshold) count -= max;
            f.skip(nbBits);
        }
        count--;  // the value read is the count + 1: -1 = "less than one"
        remaining -= count < 0 ? -count : count;
        norm[sym++] = (int16_t)count;
        if (count == 0) {  // runs of zero-probability symbols: 2-bit repeat flags, 3 = another flag follows
            for (;;) {
                const int rep = (int)f.read(2);
                for (int k = 0; k < rep; k++) {
                    if (sym > maxSym) return false;
                    norm[sym++] = 0;
                }
                if (rep != 3) break;
            }
        }
        if (remaining < 1) return false;
        while (remaining < threshold) {
            nbBits--;
            threshold >>= 1;
        }
        if (f.overrun()) return false;
    }
    if (remaining != 1 || sym > maxSym + 1 || f.overrun()) return false;
    *nsym = sym;
    *log = al;
    *used = f.bytesUsed();
    return true;
}

}  // namespace zstd
}  // namespace pgpu

#include "zstd_decode_blocks.h"
