// parquet_ingest.cpp — see parquet_ingest.h.
// Stands in for ParquetReader.nextPage / readBlock -> PrimitiveColumnReader.readPrimitive
// (lib/trino-parquet/src/main/java/io/trino/parquet/reader/ParquetReader.java:211-268, PrimitiveColumnReader.java:148-175):
// one row group's column chunks become the Blocks of one input Page. The host reads page and run headers only
// (parquet_host.h); the chunk bytes go to the device in one DMA per column and are expanded there (parquet.cu, serde.cu).
// A chunk whose data pages are all dictionary encoded becomes a DictionaryBlock (ids + the dictionary page's values, a NULL
// entry appended when the column has NULLs) — what the staged Q1 kernel consumes without materialising the values.
#include "parquet_ingest.h"

#include <algorithm>
#include <cstring>
#include <thread>

#include "parquet.h"
#include "parquet_host.h"
#include "serde.h"

namespace pgpu {

namespace {
size_t align256(size_t x) { return (x + 255) / 256 * 256; }
}  // namespace

#define CU(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            why = std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #call;        \
            return PGPU_CUDA_ERROR;                                                           \
        }                                                                                     \
    } while (0)

pgpu_status pqDecodeRowGroup(const Plan& plan, const pgpu_parquet_chunk* chunks, int32_t nchunks, int64_t rows, cudaStream_t stream, uint32_t* dErr, int pageThreads,
                             const std::function<char*(size_t)>& arena, PqDecoded& out, std::string& why) {
    if (nchunks != (int)plan.input.size())
        { why = "row group has " + std::to_string(nchunks) + " column chunks, plan expects " + std::to_string(plan.input.size()); return PGPU_INVALID_ARGUMENT; }
    if (rows < 0 || rows > 0x7fffffff) { why = "row group row count out of range"; return PGPU_INVALID_ARGUMENT; }
    std::vector<pq::Chunk> ck((size_t)nchunks);
    std::vector<std::vector<int32_t>> dictOffsets((size_t)nchunks);
    std::vector<std::vector<uint8_t>> dictBytes((size_t)nchunks);
    for (int c = 0; c < nchunks; c++) {
        const pgpu_parquet_chunk& in = chunks[c];
        const Type& t = plan.input[c];
        const int pt = in.physical_type;
        const int tl = in.type_length;
        const bool match = (pt == pq::BOOLEAN && t.kind == TypeKind::BOOLEAN) || (pt == pq::INT32 && (t.isInt32() || t.isInt64Like())) || (pt == pq::INT64 && t.isInt64Like()) ||
                           (pt == pq::DOUBLE && t.kind == TypeKind::DOUBLE) || (pt == pq::BYTE_ARRAY && t.isVarWidth()) ||
                           (pt == pq::FIXED_LEN_BYTE_ARRAY && ((tl >= 1 && tl <= 8 && t.isShortDecimal()) || (tl >= 9 && tl <= 16 && t.isLongDecimal())));
        if (!match)
            { why = "not supported on the GPU path: Parquet physical type " + std::to_string(pt) + " for a " + t.str() + " channel (channel " + std::to_string(c) + ")"; return PGPU_NOT_SUPPORTED; }
        if (in.num_values != rows) { why = "channel " + std::to_string(c) + " holds " + std::to_string(in.num_values) + " values, the row group " + std::to_string(rows) + " rows"; return PGPU_INVALID_ARGUMENT; }
    }
    // the chunks are independent and so are the pages of a chunk once their headers are known: one host thread per column chunk,
    // each sharing the remaining cores over its pages
    std::vector<std::string> whys((size_t)nchunks);
    // compressed chunks (SNAPPY / LZ4_RAW) are inflated on the host, by the chunk's thread, into a chunk with uncompressed pages
    std::vector<std::vector<uint8_t>> inflated((size_t)nchunks);
    size_t compressedTotal = 0;
    for (int c = 0; c < nchunks; c++)
        if (chunks[c].codec != pq::UNCOMPRESSED) compressedTotal += chunks[c].len;
    std::vector<const uint8_t*> cdata((size_t)nchunks);
    std::vector<size_t> clen((size_t)nchunks);
    auto parseOne = [&](int c) {
        const pgpu_parquet_chunk& in = chunks[c];
        const Type& t = plan.input[c];
        cdata[c] = in.data;
        clen[c] = in.len;
        if (in.codec != pq::UNCOMPRESSED) {
            // threads for the pages of this chunk: its share of the compressed bytes of the row group (one wide PLAIN column usually
            // holds most of them), out of the host's cores
            const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
            const int share = compressedTotal ? (int)((double)hw * (double)in.len / (double)compressedTotal + 0.5) : 1;
            if (!pq::inflateChunk(in.data, in.len, in.codec, in.num_values, inflated[c], whys[c], std::min(16, std::max(share, std::max(pageThreads, 2))))) {
                if (whys[c].empty()) whys[c] = "corrupt Parquet chunk";
                return;
            }
            cdata[c] = inflated[c].data();
            clen[c] = inflated[c].size();
        }
        if (!pq::parseChunk(cdata[c], clen[c], in.physical_type, in.max_definition_level, in.num_values, ck[c], whys[c], in.type_length, pageThreads)) {
            if (whys[c].empty()) whys[c] = "corrupt Parquet chunk";
            return;
        }
        if (in.physical_type == pq::BYTE_ARRAY && ck[c].hasDictionary) {
            const int maxChars = t.kind == TypeKind::VARCHAR ? t.precision : 0;
            if (!pq::binaryDictionary(cdata[c], ck[c].dictionaryOff, ck[c].dictionaryEnd, ck[c].dictionaryCount, maxChars, dictOffsets[c], dictBytes[c], whys[c]) && whys[c].empty())
                whys[c] = "corrupt Parquet dictionary page";
        }
    };
    if (nchunks > 1 && rows >= 65536) {
        std::vector<std::thread> workers;
        for (int c = 1; c < nchunks; c++) workers.emplace_back(parseOne, c);
        parseOne(0);
        for (auto& w : workers) w.join();
    } else {
        for (int c = 0; c < nchunks; c++) parseOne(c);
    }
    for (int c = 0; c < nchunks; c++)
        if (!whys[c].empty())
            { why = whys[c] + " (channel " + std::to_string(c) + ")"; return whys[c].rfind("not supported", 0) == 0 ? PGPU_NOT_SUPPORTED : PGPU_INVALID_ARGUMENT; }
    if (rows == 0) return PGPU_OK;  // an empty row group adds nothing
    // arena layout, every piece 256-byte aligned
    size_t need = 0;
    auto reserve = [&](size_t bytes) {
        const size_t at = need;
        need += align256(bytes + 16);
        return at;
    };
    struct Placed {
        size_t chunk = 0, levelRuns = 0, valueRuns = 0, bits = 0, nulls = 0, scratch = 0, compact = 0, values = 0, dict = 0, dictNulls = 0, dictOffsets = 0;
    };
    std::vector<Placed> at((size_t)nchunks);
    for (int c = 0; c < nchunks; c++) {
        const pq::Chunk& k = ck[c];
        Placed& q = at[c];
        const int W = plan.input[c].width();  // bytes per position of the block; 0 for var-width
        q.chunk = reserve(clen[c]);
        q.valueRuns = reserve(k.valueRuns.size() * sizeof(pq::Run));
        if (k.hasNulls()) {
            q.levelRuns = reserve(k.levelRuns.size() * sizeof(pq::Run));
            q.bits = reserve((size_t)(rows + 7) / 8);
            q.nulls = reserve((size_t)rows);
            q.scratch = reserve(serdeScratchWords(rows) * 4);
        }
        if (k.allDictionary) {
            if (k.hasNulls()) q.compact = reserve((size_t)k.nonNull * 4);
            q.values = reserve((size_t)rows * 4);  // ids
            const size_t entries = (size_t)k.dictionaryCount + 1;
            if (W) {
                q.dict = reserve(entries * W);
            } else {
                q.dictOffsets = reserve((entries + 1) * 4);
                q.dict = reserve(dictBytes[c].size());
            }
            q.dictNulls = reserve(entries);
        } else {
            if (k.hasNulls()) q.compact = reserve((size_t)k.nonNull * W);
            q.values = reserve((size_t)rows * W);
            if (k.anyDictionary) q.dict = reserve((size_t)k.dictionaryCount * W);
        }
    }
    char* base = arena(need);
    if (!base) { why = "cannot allocate " + std::to_string(need >> 20) + " MiB of device memory for the row group"; return PGPU_INSUFFICIENT_RESOURCES; }
    out.cols.assign((size_t)nchunks, pgpu_column());
    out.dicts.assign((size_t)nchunks, pgpu_column());
    std::vector<pgpu_column>&cols = out.cols, &dicts = out.dicts;
    for (int c = 0; c < nchunks; c++) {
        const pq::Chunk& k = ck[c];
        const Placed& q = at[c];
        const Type& t = plan.input[c];
        const int pt = chunks[c].physical_type;
        const int W = t.width();                                    // bytes per position of the block
        const int Win = pq::valueWidth(pt, chunks[c].type_length);  // bytes per PLAIN value in the file
        // PLAIN values that are not the block's representation: INT32 widened to 64 bits (kind 0), big-endian decimals (kind 1)
        const int conv = pt == pq::FIXED_LEN_BYTE_ARRAY ? 1 : (pt == pq::INT32 && W == 8) ? 0 : -1;
        auto plainValues = [&](size_t srcOff, int64_t count, char* dst) -> cudaError_t {
            if (count <= 0) return cudaSuccess;
            if (conv < 0) return cudaMemcpyAsync(dst, base + q.chunk + srcOff, (size_t)count * W, cudaMemcpyDeviceToDevice, stream);
            out.launches += 1;
            return pqConvertPlain(stream, (const uint8_t*)(base + q.chunk + srcOff), Win, conv, count, dst, W);
        };
        const size_t len = clen[c];
        const uint8_t* dchunk = (const uint8_t*)(base + q.chunk);
        if (len) CU(cudaMemcpyAsync(base + q.chunk, cdata[c], len, cudaMemcpyHostToDevice, stream));  // the whole chunk in one DMA
        out.h2dBytes += (int64_t)len;
        if (!k.valueRuns.empty()) CU(cudaMemcpyAsync(base + q.valueRuns, k.valueRuns.data(), k.valueRuns.size() * sizeof(pq::Run), cudaMemcpyHostToDevice, stream));
        const PqRun* vruns = (const PqRun*)(base + q.valueRuns);
        const bool nul = k.hasNulls();
        if (nul) {
            CU(cudaMemcpyAsync(base + q.levelRuns, k.levelRuns.data(), k.levelRuns.size() * sizeof(pq::Run), cudaMemcpyHostToDevice, stream));
            CU(pqLevelsToNullBits(stream, dchunk, len, (const PqRun*)(base + q.levelRuns), (int64_t)k.levelRuns.size(), rows, (uint8_t*)(base + q.bits)));
            out.launches += 1;
        }
        // where the non-null values are written: straight into the block when there is no NULL, else into `compact`
        char* dense = base + (nul ? q.compact : q.values);
        pgpu_column& col = cols[c];
        memset(&col, 0, sizeof col);
        col.type_kind = (int32_t)t.kind;
        col.precision = t.precision;
        col.scale = t.scale;
        col.position_count = rows;
        if (k.allDictionary) {
            // ids of the whole chunk in one launch
            CU(pqExpandRuns(stream, dchunk, len, vruns, (int64_t)k.valueRuns.size(), 0, k.nonNull, 4, nullptr, (uint32_t)k.dictionaryCount, dense, dErr));
            out.launches += 1;
            if (nul) {
                CU(expandNulls(stream, (const uint8_t*)(base + q.bits), rows, base + q.compact, base + q.values, (uint8_t*)(base + q.nulls), 4, (uint32_t*)(base + q.scratch)));
                CU(pqNullIds(stream, (int32_t*)(base + q.values), (const uint8_t*)(base + q.nulls), rows, (int32_t)k.dictionaryCount));
                out.launches += 4;
            }
            // the dictionary block: the dictionary page's PLAIN values + one NULL entry
            pgpu_column& d = dicts[c];
            memset(&d, 0, sizeof d);
            d.encoding = PGPU_ENC_PLAIN;
            d.type_kind = col.type_kind;
            d.precision = t.precision;
            d.scale = t.scale;
            const size_t entries = (size_t)k.dictionaryCount + 1;
            d.position_count = (int64_t)entries;
            CU(cudaMemsetAsync(base + q.dictNulls, 0, entries, stream));
            CU(cudaMemsetAsync(base + q.dictNulls + k.dictionaryCount, 1, 1, stream));
            if (nul) d.nulls = (const uint8_t*)(base + q.dictNulls);
            else d.position_count = k.dictionaryCount ? k.dictionaryCount : 1;
            if (W) {
                CU(cudaMemsetAsync(base + q.dict, 0, entries * W, stream));
                CU(plainValues(k.dictionaryOff, k.dictionaryCount, base + q.dict));
                d.values = base + q.dict;
            } else {
                std::vector<int32_t>& off = dictOffsets[c];
                off.push_back(off.back());  // the NULL entry is empty
                CU(cudaMemcpyAsync(base + q.dictOffsets, off.data(), off.size() * 4, cudaMemcpyHostToDevice, stream));
                if (!dictBytes[c].empty()) CU(cudaMemcpyAsync(base + q.dict, dictBytes[c].data(), dictBytes[c].size(), cudaMemcpyHostToDevice, stream));
                d.offsets = (const int32_t*)(base + q.dictOffsets);
                d.values = base + q.dict;
            }
            col.encoding = PGPU_ENC_DICTIONARY;
            col.ids = (const int32_t*)(base + q.values);
            col.dictionary = &d;
            continue;
        }
        col.encoding = PGPU_ENC_PLAIN;
        col.values = base + q.values;
        if (pt == pq::BOOLEAN) {
            CU(pqExpandRuns(stream, dchunk, len, vruns, (int64_t)k.valueRuns.size(), 0, k.nonNull, 1, nullptr, 2u, dense, dErr));
            out.launches += 1;
        } else {
            if (k.anyDictionary) CU(plainValues(k.dictionaryOff, k.dictionaryCount, base + q.dict));
            for (const pq::DataPage& pg : k.pages) {
                if (!pg.nonNull) continue;
                char* dst = dense + (size_t)pg.firstValue * W;
                if (pg.dictionaryEncoded) {  // a chunk that fell back to PLAIN part-way: materialise the dictionary-encoded pages
                    CU(pqExpandRuns(stream, dchunk, len, vruns + pg.runLo, (int64_t)(pg.runHi - pg.runLo), pg.firstValue, pg.nonNull, W, base + q.dict,
                                    (uint32_t)k.dictionaryCount, dst, dErr));
                    out.launches += 1;
                } else {
                    CU(plainValues(pg.valuesOff, pg.nonNull, dst));
                }
            }
        }
        if (nul) {
            col.nulls = (const uint8_t*)(base + q.nulls);
            CU(expandNulls(stream, (const uint8_t*)(base + q.bits), rows, base + q.compact, base + q.values, (uint8_t*)(base + q.nulls), W, (uint32_t*)(base + q.scratch)));
            out.launches += 3;
        }
    }
    return PGPU_OK;
}

#undef CU

}  // namespace pgpu
