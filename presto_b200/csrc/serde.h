// serde.h — device side of ingesting a page in the reference's wire format (main/execution/buffer/PagesSerdeUtil.java:45-62,
// spi/block/*BlockEncoding.java): the serialized payload is copied to the device as it is (one DMA) and decoded there.
// What needs work on the device is the null handling: flags travel as a bit stream, most significant bit first
// (spi/block/EncoderUtil.java:33-70), and a fixed-width block that may have nulls carries only its non-null values
// (spi/block/LongArrayBlockEncoding.java:45-60), so positions have to be re-expanded with a prefix sum over the flags.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pgpu {

constexpr int kSerdeBlockPositions = 1024;  // positions per thread block (256 threads x 4)

// number of uint32 scratch words expandNulls needs for n positions
inline size_t serdeScratchWords(int64_t n) { return (size_t)((n + kSerdeBlockPositions - 1) / kSerdeBlockPositions) + 1; }

// bits -> nulls[n] (1 = NULL); compact[nonNull] (aligned to `width`) -> values[n], zero at NULL positions.
// width 1 / 4 / 8 / 16; `scratch` = serdeScratchWords(n) words of device memory. Three launches on `stream`.
cudaError_t expandNulls(cudaStream_t stream, const uint8_t* bits, int64_t n, const void* compact, void* values, uint8_t* nulls, int width, uint32_t* scratch);

// bits -> nulls[n] only (var-width blocks keep zero-length entries for NULL positions)
cudaError_t unpackNullBits(cudaStream_t stream, const uint8_t* bits, int64_t n, uint8_t* nulls);

// ---- the other direction: output blocks -> wire pieces (getOutput in the wire format) ----------------------------------------
constexpr int kPackBlockPositions = 2048;  // positions per thread block (256 threads x 8 = one byte of null bits per thread)

inline size_t packScratchWords(int64_t n) { return (size_t)((n + kPackBlockPositions - 1) / kPackBlockPositions) + 1; }

// nulls[n] (1 = NULL) -> bits[(n + 7) / 8], most significant bit first; counts[block] = non-null positions of the block, then
// scanned in place (exclusive), counts[blocks] = total non-null. Two launches.
cudaError_t packNullBits(cudaStream_t stream, const uint8_t* nulls, int64_t n, uint8_t* bits, uint32_t* counts);

// values[n] -> compact[non-null], in position order; `counts` as left by packNullBits. width 1 / 4 / 8 / 16. One launch.
cudaError_t compactNonNull(cudaStream_t stream, const uint8_t* nulls, int64_t n, const void* values, void* compact, int width, const uint32_t* counts);

// fixed-stride cells (lenPrefixed: byte 0 = length, data from byte 1; else `stride` data bytes; NULL = empty) ->
// counts (byte totals per block, scanned; counts[blocks] = total bytes). Two launches.
cudaError_t cellByteCounts(cudaStream_t stream, const uint8_t* cells, const uint8_t* nulls, int64_t n, int stride, bool lenPrefixed, uint32_t* counts);

// ... -> endOffsets[n] (int32, the wire's END offsets) and the concatenated bytes. One launch.
cudaError_t packCells(cudaStream_t stream, const uint8_t* cells, const uint8_t* nulls, int64_t n, int stride, bool lenPrefixed, const uint32_t* counts,
                      int32_t* endOffsets, uint8_t* bytes);

}  // namespace pgpu
