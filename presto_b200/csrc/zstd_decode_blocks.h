// zstd_decode_blocks.h — second half of zstd_decode.h: Huffman literals, sequences, blocks, frames.
#pragma once

namespace pgpu {
namespace zstd {

// ---- Huffman (RFC 8878 4.2): at most 256 symbols, codes of at most 11 bits
struct HufTable {
    int log = 0;  // 0 = no table yet
    std::vector<uint8_t> symbol, bits;
};

// weights[0 .. n) of the first n symbols; the last symbol's weight completes the sum of 2^(w - 1) to a power of two
inline bool hufBuild(HufTable& t, uint8_t* weights, int n) {
    if (n < 1 || n > 255) return false;
    uint32_t total = 0;
    for (int s = 0; s < n; s++) {
        if (weights[s] > 11) return false;
        if (weights[s]) total += 1u << (weights[s] - 1);
    }
    if (!total) return false;
    const int log = highBit(total) + 1;
    if (log > 11) return false;
    const uint32_t rest = (1u << log) - total;
    if (rest & (rest - 1)) return false;  // must be a power of two
    weights[n] = (uint8_t)(highBit(rest) + 1);
    const int nsym = n + 1;
    uint32_t rankCount[13] = {0}, rankStart[13] = {0};
    for (int s = 0; s < nsym; s++) rankCount[weights[s]]++;
    if (rankCount[1] < 2 || (rankCount[1] & 1)) return false;  // a complete code has an even number (>= 2) of longest codes
    uint32_t at = 0;
    for (int w = 1; w <= log; w++) {
        rankStart[w] = at;
        at += rankCount[w] << (w - 1);
    }
    t.log = log;
    t.symbol.assign((size_t)1 << log, 0);
    t.bits.assign((size_t)1 << log, 0);
    for (int s = 0; s < nsym; s++) {
        const int w = weights[s];
        if (!w) continue;
        const uint32_t len = 1u << (w - 1);
        for (uint32_t k = 0; k < len; k++) {
            t.symbol[rankStart[w] + k] = (uint8_t)s;
            t.bits[rankStart[w] + k] = (uint8_t)(log + 1 - w);
        }
        rankStart[w] += len;
    }
    return true;
}

// the tree description in front of the literal streams; *used = its bytes
inline bool hufReadTable(HufTable& t, const uint8_t* d, size_t n, size_t* used) {
    if (!n) return false;
    uint8_t weights[256] = {0};
    const int hb = d[0];
    int count = 0;
    if (hb >= 128) {  // direct: 4 bits per weight, high nibble first
        count = hb - 127;
        const size_t bytes = ((size_t)count + 1) / 2;
        if (1 + bytes > n) return false;
        for (int k = 0; k < count; k++) weights[k] = (k & 1) ? (uint8_t)(d[1 + k / 2] & 15) : (uint8_t)(d[1 + k / 2] >> 4);
        *used = 1 + bytes;
    } else {  // FSE-compressed weights: table description, then two interleaved states over a backward bitstream
        if (!hb || 1 + (size_t)hb > n) return false;
        int16_t norm[16];
        int nsym = 0, log = 0;
        size_t head = 0;
        if (!fseReadCounts(d + 1, (size_t)hb, 6, 11, norm, &nsym, &log, &head)) return false;
        FseTable ft;
        if (!fseBuild(ft, norm, nsym, log)) return false;
        Backward b;
        if (head >= (size_t)hb || !b.init(d + 1 + head, (size_t)hb - head)) return false;
        uint32_t s1 = b.read(log), s2 = b.read(log);
        if (b.pos < 0) return false;
        for (;;) {
            if (count > 253) return false;
            weights[count++] = ft.cell[s1].symbol;
            s1 = ft.cell[s1].newState + b.read(ft.cell[s1].bits);
            if (b.pos < 0) {
                weights[count++] = ft.cell[s2].symbol;
                break;
            }
            if (count > 253) return false;
            weights[count++] = ft.cell[s2].symbol;
            s2 = ft.cell[s2].newState + b.read(ft.cell[s2].bits);
            if (b.pos < 0) {
                weights[count++] = ft.cell[s1].symbol;
                break;
            }
        }
        *used = 1 + (size_t)hb;
    }
    return hufBuild(t, weights, count);
}

// one backward stream -> exactly `count` symbols; the stream must end exactly where its first byte begins
inline bool hufDecodeStream(const HufTable& t, const uint8_t* d, size_t n, uint8_t* out, size_t count) {
    Backward b;
    if (!b.init(d, n)) return false;
    for (size_t k = 0; k < count; k++) {
        const uint32_t idx = b.peek(t.log);  // bits past the start read as zero
        out[k] = t.symbol[idx];
        b.pos -= t.bits[idx];
        if (b.pos < 0) return false;
    }
    return b.pos == 0;
}

// ---- sequences
struct SeqTables {
    FseTable ll, of, ml;
    bool have = false;  // tables of an earlier block exist (Repeat_Mode)
};

static const int16_t kLLDefault[36] = {4, 3, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 3, 2, 1, 1, 1, 1, 1, -1, -1, -1, -1};
static const int16_t kMLDefault[53] = {1, 4, 3, 2, 2, 2, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1,
                                       1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1};
static const int16_t kOFDefault[29] = {1, 1, 1, 1, 1, 1, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1};
static const uint32_t kLLBase[36] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 22, 24, 28, 32, 40, 48, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 65536};
static const uint8_t kLLBits[36] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 3, 3, 4, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};
static const uint32_t kMLBase[53] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 27, 28, 29, 30, 31, 32, 33, 34, 35, 37, 39, 41,
                                     43, 47, 51, 59, 67, 83, 99, 131, 259, 515, 1027, 2051, 4099, 8195, 16387, 32771, 65539};
static const uint8_t kMLBits[53] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1,
                                    2, 2, 3, 3, 4, 4, 5, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};

// one of the three code tables of a sequences section: mode 0 predefined, 1 RLE, 2 described, 3 the previous block's
inline bool seqTable(FseTable& t, int mode, const uint8_t* d, size_t n, size_t* used, const int16_t* dflt, int dfltSyms, int dfltLog, int maxLog, int maxSym, bool havePrev) {
    *used = 0;
    if (mode == 0) return fseBuild(t, dflt, dfltSyms, dfltLog);
    if (mode == 1) {
        if (!n || d[0] > maxSym) return false;
        fseRle(t, d[0]);
        *used = 1;
        return true;
    }
    if (mode == 2) {
        int16_t norm[64];
        int nsym = 0, log = 0;
        if (!fseReadCounts(d, n, maxLog, maxSym, norm, &nsym, &log, used)) return false;
        return fseBuild(t, norm, nsym, log);
    }
    return havePrev && !t.cell.empty();
}

struct BlockState {
    HufTable huf;
    SeqTables seq;
    uint32_t rep[3] = {1, 4, 8};
};

// one Compressed_Block: literals section + sequences section -> appended to out[op ..)
inline bool decodeCompressedBlock(BlockState& st, const uint8_t* d, size_t n, uint8_t* out, size_t outLen, size_t& op, std::vector<uint8_t>& lit) {
    if (n < 2) return false;
    // ---- literals
    const int ltype = d[0] & 3, sf = (d[0] >> 2) & 3;
    size_t regen = 0, comp = 0, head = 0;
    bool four = false;
    if (ltype < 2) {
        if (sf == 0 || sf == 2) { regen = d[0] >> 3; head = 1; }
        else if (sf == 1) { regen = (size_t)(d[0] >> 4) | ((size_t)d[1] << 4); head = 2; }
        else {
            if (n < 3) return false;
            regen = (size_t)(d[0] >> 4) | ((size_t)d[1] << 4) | ((size_t)d[2] << 12);
            head = 3;
        }
    } else {
        if (n < 5) return false;
        const uint64_t h = (uint64_t)d[0] | ((uint64_t)d[1] << 8) | ((uint64_t)d[2] << 16) | ((uint64_t)d[3] << 24) | ((uint64_t)d[4] << 32);
        four = sf != 0;
        if (sf < 2) { regen = (h >> 4) & 0x3ff; comp = (h >> 14) & 0x3ff; head = 3; }
        else if (sf == 2) { regen = (h >> 4) & 0x3fff; comp = (h >> 18) & 0x3fff; head = 4; }
        else { regen = (h >> 4) & 0x3ffff; comp = (h >> 22) & 0x3ffff; head = 5; }
    }
    if (regen > (128u << 10)) return false;  // Block_Maximum_Size
    lit.assign(regen + 8, 0);
    size_t at = head;
    if (ltype == 0) {
        if (regen > n - at) return false;
        memcpy(lit.data(), d + at, regen);
        at += regen;
    } else if (ltype == 1) {
        if (at >= n) return false;
        memset(lit.data(), d[at], regen);
        at += 1;
    } else {
        if (comp > n - at) return false;
        const uint8_t* p = d + at;
        size_t left = comp;
        if (ltype == 2) {
            size_t used = 0;
            if (!hufReadTable(st.huf, p, left, &used)) return false;
            p += used;
            left -= used;
        } else if (!st.huf.log) {
            return false;  // Treeless_Literals_Block without an earlier tree
        }
        if (!four) {
            if (!hufDecodeStream(st.huf, p, left, lit.data(), regen)) return false;
        } else {
            if (left < 6) return false;
            const size_t s1 = (size_t)p[0] | ((size_t)p[1] << 8), s2 = (size_t)p[2] | ((size_t)p[3] << 8), s3 = (size_t)p[4] | ((size_t)p[5] << 8);
            if (s1 + s2 + s3 > left - 6) return false;
            const size_t s4 = left - 6 - s1 - s2 - s3, per = (regen + 3) / 4;
            if (3 * per > regen) return false;
            const uint8_t* q = p + 6;
            if (!hufDecodeStream(st.huf, q, s1, lit.data(), per) || !hufDecodeStream(st.huf, q + s1, s2, lit.data() + per, per) ||
                !hufDecodeStream(st.huf, q + s1 + s2, s3, lit.data() + 2 * per, per) || !hufDecodeStream(st.huf, q + s1 + s2 + s3, s4, lit.data() + 3 * per, regen - 3 * per))
                return false;
        }
        at += comp;
    }
    // ---- sequences
    if (at >= n) return false;
    size_t nseq = d[at++];
    if (nseq >= 128) {
        if (nseq == 255) {
            if (at + 2 > n) return false;
            nseq = (size_t)d[at] + ((size_t)d[at + 1] << 8) + 0x7f00;
            at += 2;
        } else {
            if (at >= n) return false;
            nseq = ((nseq - 128) << 8) + d[at++];
        }
    }
    size_t litAt = 0;
    if (nseq) {
        if (at >= n) return false;
        const int modes = d[at++];
        if (modes & 3) return false;  // reserved bits
        size_t used = 0;
        if (!seqTable(st.seq.ll, modes >> 6, d + at, n - at, &used, kLLDefault, 36, 6, 9, 35, st.seq.have)) return false;
        at += used;
        if (!seqTable(st.seq.of, (modes >> 4) & 3, d + at, n - at, &used, kOFDefault, 29, 5, 8, 31, st.seq.have)) return false;
        at += used;
        if (!seqTable(st.seq.ml, (modes >> 2) & 3, d + at, n - at, &used, kMLDefault, 53, 6, 9, 52, st.seq.have)) return false;
        at += used;
        st.seq.have = true;
        Backward b;
        if (at >= n || !b.init(d + at, n - at)) return false;
        uint32_t sll = b.read(st.seq.ll.log), sof = b.read(st.seq.of.log), sml = b.read(st.seq.ml.log);
        if (b.pos < 0) return false;
        for (size_t k = 0; k < nseq; k++) {
            const int ofCode = st.seq.of.cell[sof].symbol, mlCode = st.seq.ml.cell[sml].symbol, llCode = st.seq.ll.cell[sll].symbol;
            if (ofCode > 31 || mlCode > 52 || llCode > 35) return false;
            const uint64_t ofValue = ((uint64_t)1 << ofCode) + b.read(ofCode);  // (ofCode <= 31: one read of up to 31 bits)
            const uint32_t ml = kMLBase[mlCode] + b.read(kMLBits[mlCode]);
            const uint32_t ll = kLLBase[llCode] + b.read(kLLBits[llCode]);
            if (b.pos < 0) return false;
            uint32_t offset;
            if (ofValue > 3) {
                offset = (uint32_t)(ofValue - 3);
                st.rep[2] = st.rep[1];
                st.rep[1] = st.rep[0];
                st.rep[0] = offset;
            } else {
                uint32_t idx = (uint32_t)ofValue - 1 + (ll == 0 ? 1u : 0u);  // 0, 1, 2, or 3 = "first repeat offset minus one"
                if (idx == 0) {
                    offset = st.rep[0];
                } else {
                    offset = idx < 3 ? st.rep[idx] : st.rep[0] - 1;
                    if (!offset) return false;
                    if (idx > 1) st.rep[2] = st.rep[1];
                    st.rep[1] = st.rep[0];
                    st.rep[0] = offset;
                }
            }
            if (k + 1 < nseq) {  // states move on after every sequence but the last: literal length, match length, offset
                sll = st.seq.ll.cell[sll].newState + b.read(st.seq.ll.cell[sll].bits);
                sml = st.seq.ml.cell[sml].newState + b.read(st.seq.ml.cell[sml].bits);
                sof = st.seq.of.cell[sof].newState + b.read(st.seq.of.cell[sof].bits);
                if (b.pos < 0) return false;
            }
            if (ll > regen - litAt || (size_t)ll + ml > outLen - op) return false;
            memcpy(out + op, lit.data() + litAt, ll);
            litAt += ll;
            op += ll;
            if (offset > op) return false;  // no dictionary: a match cannot start before the frame's first byte
            if (offset >= ml) memcpy(out + op, out + op - offset, ml);
            else for (uint32_t j = 0; j < ml; j++) out[op + j] = out[op - offset + j];
            op += ml;
        }
        if (b.pos != 0) return false;
    }
    const size_t tail = regen - litAt;
    if (tail > outLen - op) return false;
    memcpy(out + op, lit.data() + litAt, tail);
    op += tail;
    return true;
}

// one or more frames (skippable frames are skipped) -> exactly outLen bytes
inline bool decompress(const uint8_t* in, size_t n, uint8_t* out, size_t outLen) {
    size_t ip = 0, op = 0;
    std::vector<uint8_t> lit;
    while (ip < n) {
        if (n - ip < 4) return false;
        const uint32_t magic = (uint32_t)in[ip] | ((uint32_t)in[ip + 1] << 8) | ((uint32_t)in[ip + 2] << 16) | ((uint32_t)in[ip + 3] << 24);
        ip += 4;
        if ((magic & 0xfffffff0u) == 0x184d2a50u) {  // skippable frame
            if (n - ip < 4) return false;
            const size_t len = (size_t)in[ip] | ((size_t)in[ip + 1] << 8) | ((size_t)in[ip + 2] << 16) | ((size_t)in[ip + 3] << 24);
            ip += 4;
            if (len > n - ip) return false;
            ip += len;
            continue;
        }
        if (magic != 0xfd2fb528u || ip >= n) return false;
        const uint8_t fhd = in[ip++];
        if (fhd & 8) return false;  // reserved bit
        const bool single = (fhd & 32) != 0, checksum = (fhd & 4) != 0;
        const int dictBytes = (fhd & 3) == 3 ? 4 : (fhd & 3), fcsFlag = fhd >> 6;
        const int fcsBytes = fcsFlag == 0 ? (single ? 1 : 0) : fcsFlag == 1 ? 2 : fcsFlag == 2 ? 4 : 8;
        const size_t headLeft = (single ? 0 : 1) + (size_t)dictBytes + (size_t)fcsBytes;
        if (headLeft > n - ip) return false;
        if (!single) ip++;  // window descriptor: matches are bounded by what has been produced, which is stricter
        for (int k = 0; k < dictBytes; k++)
            if (in[ip + k]) return false;  // a dictionary id: Parquet pages carry no dictionary
        ip += (size_t)dictBytes;
        uint64_t fcs = 0;
        for (int k = 0; k < fcsBytes; k++) fcs |= (uint64_t)in[ip + k] << (8 * k);
        if (fcsBytes == 2) fcs += 256;
        ip += (size_t)fcsBytes;
        const size_t frameStart = op;
        BlockState st;
        for (;;) {
            if (n - ip < 3) return false;
            const uint32_t bh = (uint32_t)in[ip] | ((uint32_t)in[ip + 1] << 8) | ((uint32_t)in[ip + 2] << 16);
            ip += 3;
            const bool last = bh & 1;
            const int type = (bh >> 1) & 3;
            const size_t size = bh >> 3;
            if (type == 0) {
                if (size > n - ip || size > outLen - op) return false;
                memcpy(out + op, in + ip, size);
                ip += size;
                op += size;
            } else if (type == 1) {
                if (ip >= n || size > outLen - op) return false;
                memset(out + op, in[ip], size);
                ip += 1;
                op += size;
            } else if (type == 2) {
                if (size > n - ip || size > (128u << 10)) return false;
                // matches may reach back into earlier blocks of this frame, not before it
                size_t rel = op - frameStart;
                if (!decodeCompressedBlock(st, in + ip, size, out + frameStart, outLen - frameStart, rel, lit)) return false;
                op = frameStart + rel;
                ip += size;
            } else {
                return false;
            }
            if (last) break;
        }
        if (fcsBytes && fcs != (uint64_t)(op - frameStart)) return false;
        if (checksum) {
            if (n - ip < 4) return false;
            ip += 4;  // XXH64 of the content, low 32 bits: not verified (the declared sizes are)
        }
    }
    return op == outLen;
}

}  // namespace zstd
}  // namespace pgpu
