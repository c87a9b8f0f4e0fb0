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

}  // namespace pgpu
