// parquet.h — device side of ingesting Parquet column chunks (see parquet_host.h for what the host reads): the chunk's bytes
// are in HBM as they lie in the file; these kernels expand the RLE / bit-packed hybrid streams (definition levels, dictionary
// ids, booleans) that the reference decodes value by value in PrimitiveColumnReader.readValues
// (lib/trino-parquet/src/main/java/io/trino/parquet/reader/PrimitiveColumnReader.java:185-225) through parquet-mr's
// RunLengthBitPackingHybridDecoder. One thread per output value: the run is found by binary search in the run table the host
// built from the run headers, a bit-packed value is extracted from at most 5 bytes. HBM-bound byte work, no tensor cores.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pgpu {

struct PqRun {  // = pq::Run (parquet_host.h)
    int64_t start;
    uint64_t src;
    uint32_t count;
    uint32_t widthPacked;
};

constexpr uint32_t kPqErrBadId = 64u;  // PGPU_ERR_PARQUET_ID: a dictionary id >= the dictionary size (corrupt page)

// definition levels (max level 1) of rows [0, n) -> null flags in the page wire format's bit order (most significant bit
// first, 1 = NULL), ready for expandNulls (serde.h). `runs` must cover [0, n) without gaps.
cudaError_t pqLevelsToNullBits(cudaStream_t stream, const uint8_t* chunk, size_t chunkLen, const PqRun* runs, int64_t nruns, int64_t n, uint8_t* bits);

// values [first, first + count) of a hybrid stream -> out[0 .. count). outWidth 4 with dict == nullptr: the codes themselves
// (dictionary ids; a code >= limit raises kPqErrBadId in *err and is stored as 0). dict != nullptr: out = dict[code], outWidth 4
// 8 or 16 bytes per value. outWidth 1: the codes as bytes (BOOLEAN).
cudaError_t pqExpandRuns(cudaStream_t stream, const uint8_t* chunk, size_t chunkLen, const PqRun* runs, int64_t nruns, int64_t first, int64_t count,
                         int outWidth, const void* dict, uint32_t limit, void* out, uint32_t* err);

// PLAIN values as they lie in the chunk (any alignment) -> block values. kind 0: little-endian int32 -> int64 (IntColumnReader /
// ShortDecimalColumnReader on an INT32 column). kind 1: big-endian two's complement of inWidth (1..16) bytes -> int64 (outWidth 8,
// inWidth <= 8: ParquetTypeUtils.getShortDecimalValue) or -> the 128-bit sign-magnitude form of Int128ArrayBlock, low word first,
// sign in the top bit of the high word (outWidth 16: Decimals.encodeUnscaledValue(new BigInteger(bytes))).
cudaError_t pqConvertPlain(cudaStream_t stream, const uint8_t* src, int inWidth, int kind, int64_t count, void* out, int outWidth);

// ids[i] = nullId where nulls[i] (the NULL entry appended to the dictionary)
cudaError_t pqNullIds(cudaStream_t stream, int32_t* ids, const uint8_t* nulls, int64_t n, int32_t nullId);

}  // namespace pgpu
