// serde.cu — see serde.h. Ahead-of-time sm_100a kernels.
#include "serde.h"

namespace pgpu {
namespace {

constexpr int kThreads = 256;
constexpr unsigned kFull = 0xffffffffu;

__device__ inline bool isNullBit(const uint8_t* bits, int64_t p) { return (bits[p >> 3] >> (7 - (int)(p & 7))) & 1; }

// the 4 positions of a thread: bit i of the result = position base + i is NOT null (positions >= n count as null)
__device__ inline unsigned liveMask(const uint8_t* bits, int64_t base, int64_t n) {
    unsigned m = 0;
#pragma unroll
    for (int i = 0; i < 4; i++)
        if (base + i < n && !isNullBit(bits, base + i)) m |= 1u << i;
    return m;
}

__global__ void __launch_bounds__(kThreads) k_live_counts(const uint8_t* bits, int64_t n, uint32_t* counts) {
    __shared__ unsigned s_total;
    if (threadIdx.x == 0) s_total = 0;
    __syncthreads();
    const int64_t base = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 4;
    unsigned c = __popc(liveMask(bits, base, n));
    c = __reduce_add_sync(kFull, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_total, c);
    __syncthreads();
    if (threadIdx.x == 0) counts[blockIdx.x] = s_total;
}

// exclusive scan of counts[0 .. m) in place, counts[m] = total; one block
__global__ void __launch_bounds__(1024) k_scan_counts(uint32_t* counts, int64_t m) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t i0 = 0; i0 < m; i0 += 1024) {
        const int64_t i = i0 + threadIdx.x;
        const uint32_t v = i < m ? counts[i] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(kFull, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const uint32_t w = s_warp[lane];
            uint32_t wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(kFull, wi, d);
                if (lane >= d) wi += t;
            }
            s_warp[lane] = wi - w;
        }
        __syncthreads();
        const uint32_t carry = s_carry;
        if (i < m) counts[i] = carry + s_warp[warp] + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + s_warp[warp] + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) counts[m] = s_carry;
}

template <typename T>
__device__ inline T zeroOf() { return T(0); }
template <>
__device__ inline uint4 zeroOf<uint4>() { return make_uint4(0, 0, 0, 0); }

template <typename T>
__global__ void __launch_bounds__(kThreads) k_expand(const uint8_t* bits, int64_t n, const T* compact, T* values, uint8_t* nulls, const uint32_t* blockBase) {
    __shared__ uint32_t s_warp[kThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t base = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 4;
    const unsigned m = liveMask(bits, base, n);
    const uint32_t cnt = __popc(m);
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(kFull, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t before = 0;
    for (int w = 0; w < warp; w++) before += s_warp[w];
    uint64_t src = (uint64_t)blockBase[blockIdx.x] + before + incl - cnt;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (base + i >= n) break;
        const bool live = (m >> i) & 1;
        values[base + i] = live ? compact[src] : zeroOf<T>();
        nulls[base + i] = live ? 0 : 1;
        src += live ? 1 : 0;
    }
}

__global__ void __launch_bounds__(kThreads) k_unpack_bits(const uint8_t* bits, int64_t n, uint8_t* nulls) {
    for (int64_t p = (int64_t)blockIdx.x * kThreads + threadIdx.x; p < n; p += (int64_t)gridDim.x * kThreads) nulls[p] = isNullBit(bits, p) ? 1 : 0;
}

// ---- encode ------------------------------------------------------------------------------------------------------------
constexpr int kPer = 8;  // positions per thread

// block-wide exclusive prefix of `v` (256 threads); s_warp: 8 words
__device__ inline uint32_t blockExclusive(uint32_t v, uint32_t* s_warp) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(kFull, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t before = 0;
    for (int w = 0; w < warp; w++) before += s_warp[w];
    return before + incl - v;
}

__device__ inline void blockTotal(uint32_t v, uint32_t* out) {
    __shared__ unsigned s_total;
    if (threadIdx.x == 0) s_total = 0;
    __syncthreads();
    v = __reduce_add_sync(kFull, v);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(&s_total, v);
    __syncthreads();
    if (threadIdx.x == 0) *out = s_total;
}

__global__ void __launch_bounds__(kThreads) k_pack_nulls(const uint8_t* nulls, int64_t n, uint8_t* bits, uint32_t* counts) {
    const int64_t base = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * kPer;
    unsigned byte = 0, live = 0;
#pragma unroll
    for (int i = 0; i < kPer; i++)
        if (base + i < n) {
            const bool isNull = nulls[base + i] != 0;
            byte |= (isNull ? 1u : 0u) << (7 - i);
            live += isNull ? 0u : 1u;
        }
    if (base < n) bits[base >> 3] = (uint8_t)byte;
    blockTotal(live, counts + blockIdx.x);
}

template <typename T>
__global__ void __launch_bounds__(kThreads) k_compact(const uint8_t* nulls, int64_t n, const T* values, T* compact, const uint32_t* blockBase) {
    __shared__ uint32_t s_warp[kThreads / 32];
    const int64_t base = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * kPer;
    unsigned m = 0;
#pragma unroll
    for (int i = 0; i < kPer; i++)
        if (base + i < n && nulls[base + i] == 0) m |= 1u << i;
    uint64_t dst = (uint64_t)blockBase[blockIdx.x] + blockExclusive(__popc(m), s_warp);
#pragma unroll
    for (int i = 0; i < kPer; i++)
        if ((m >> i) & 1) compact[dst++] = values[base + i];
}

__device__ inline uint32_t cellLength(const uint8_t* cells, const uint8_t* nulls, int64_t p, int stride, bool lenPrefixed) {
    if (nulls[p]) return 0;
    return lenPrefixed ? cells[(size_t)p * stride] : (uint32_t)stride;
}

__global__ void __launch_bounds__(kThreads) k_cell_bytes(const uint8_t* cells, const uint8_t* nulls, int64_t n, int stride, bool lenPrefixed, uint32_t* counts) {
    const int64_t base = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * kPer;
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < kPer; i++)
        if (base + i < n) sum += cellLength(cells, nulls, base + i, stride, lenPrefixed);
    blockTotal(sum, counts + blockIdx.x);
}

__global__ void __launch_bounds__(kThreads) k_pack_cells(const uint8_t* cells, const uint8_t* nulls, int64_t n, int stride, bool lenPrefixed, const uint32_t* blockBase,
                                                         int32_t* endOffsets, uint8_t* bytes) {
    __shared__ uint32_t s_warp[kThreads / 32];
    const int64_t base = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * kPer;
    uint32_t len[kPer], sum = 0;
#pragma unroll
    for (int i = 0; i < kPer; i++) {
        len[i] = base + i < n ? cellLength(cells, nulls, base + i, stride, lenPrefixed) : 0u;
        sum += len[i];
    }
    uint32_t at = blockBase[blockIdx.x] + blockExclusive(sum, s_warp);
#pragma unroll
    for (int i = 0; i < kPer; i++) {
        if (base + i >= n) break;
        const uint8_t* src = cells + (size_t)(base + i) * stride + (lenPrefixed ? 1 : 0);
        for (uint32_t k = 0; k < len[i]; k++) bytes[at + k] = src[k];
        at += len[i];
        endOffsets[base + i] = (int32_t)at;
    }
}

}  // namespace

cudaError_t packNullBits(cudaStream_t stream, const uint8_t* nulls, int64_t n, uint8_t* bits, uint32_t* counts) {
    if (n <= 0) return cudaSuccess;
    const int64_t nblocks = (n + kPackBlockPositions - 1) / kPackBlockPositions;
    k_pack_nulls<<<(unsigned)nblocks, kThreads, 0, stream>>>(nulls, n, bits, counts);
    k_scan_counts<<<1, 1024, 0, stream>>>(counts, nblocks);
    return cudaGetLastError();
}

cudaError_t compactNonNull(cudaStream_t stream, const uint8_t* nulls, int64_t n, const void* values, void* compact, int width, const uint32_t* counts) {
    if (n <= 0) return cudaSuccess;
    const unsigned nblocks = (unsigned)((n + kPackBlockPositions - 1) / kPackBlockPositions);
    switch (width) {
        case 1: k_compact<uint8_t><<<nblocks, kThreads, 0, stream>>>(nulls, n, (const uint8_t*)values, (uint8_t*)compact, counts); break;
        case 4: k_compact<uint32_t><<<nblocks, kThreads, 0, stream>>>(nulls, n, (const uint32_t*)values, (uint32_t*)compact, counts); break;
        case 8: k_compact<uint64_t><<<nblocks, kThreads, 0, stream>>>(nulls, n, (const uint64_t*)values, (uint64_t*)compact, counts); break;
        case 16: k_compact<uint4><<<nblocks, kThreads, 0, stream>>>(nulls, n, (const uint4*)values, (uint4*)compact, counts); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t cellByteCounts(cudaStream_t stream, const uint8_t* cells, const uint8_t* nulls, int64_t n, int stride, bool lenPrefixed, uint32_t* counts) {
    if (n <= 0) return cudaSuccess;
    const int64_t nblocks = (n + kPackBlockPositions - 1) / kPackBlockPositions;
    k_cell_bytes<<<(unsigned)nblocks, kThreads, 0, stream>>>(cells, nulls, n, stride, lenPrefixed, counts);
    k_scan_counts<<<1, 1024, 0, stream>>>(counts, nblocks);
    return cudaGetLastError();
}

cudaError_t packCells(cudaStream_t stream, const uint8_t* cells, const uint8_t* nulls, int64_t n, int stride, bool lenPrefixed, const uint32_t* counts,
                      int32_t* endOffsets, uint8_t* bytes) {
    if (n <= 0) return cudaSuccess;
    const unsigned nblocks = (unsigned)((n + kPackBlockPositions - 1) / kPackBlockPositions);
    k_pack_cells<<<nblocks, kThreads, 0, stream>>>(cells, nulls, n, stride, lenPrefixed, counts, endOffsets, bytes);
    return cudaGetLastError();
}

cudaError_t expandNulls(cudaStream_t stream, const uint8_t* bits, int64_t n, const void* compact, void* values, uint8_t* nulls, int width, uint32_t* scratch) {
    if (n <= 0) return cudaSuccess;
    const int64_t nblocks = (n + kSerdeBlockPositions - 1) / kSerdeBlockPositions;
    k_live_counts<<<(unsigned)nblocks, kThreads, 0, stream>>>(bits, n, scratch);
    k_scan_counts<<<1, 1024, 0, stream>>>(scratch, nblocks);
    switch (width) {
        case 1: k_expand<uint8_t><<<(unsigned)nblocks, kThreads, 0, stream>>>(bits, n, (const uint8_t*)compact, (uint8_t*)values, nulls, scratch); break;
        case 4: k_expand<uint32_t><<<(unsigned)nblocks, kThreads, 0, stream>>>(bits, n, (const uint32_t*)compact, (uint32_t*)values, nulls, scratch); break;
        case 8: k_expand<uint64_t><<<(unsigned)nblocks, kThreads, 0, stream>>>(bits, n, (const uint64_t*)compact, (uint64_t*)values, nulls, scratch); break;
        case 16: k_expand<uint4><<<(unsigned)nblocks, kThreads, 0, stream>>>(bits, n, (const uint4*)compact, (uint4*)values, nulls, scratch); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t unpackNullBits(cudaStream_t stream, const uint8_t* bits, int64_t n, uint8_t* nulls) {
    if (n <= 0) return cudaSuccess;
    int64_t g = (n + kThreads - 1) / kThreads;
    if (g > 148 * 16) g = 148 * 16;
    k_unpack_bits<<<(unsigned)g, kThreads, 0, stream>>>(bits, n, nulls);
    return cudaGetLastError();
}

}  // namespace pgpu

// ---- pgpu_expand_nulls_static: expandNulls with its scratch in a file-scope device array (no allocation) --------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it). The three launches are those of expandNulls above;
// the per-block counts live in g_expandScratch instead of a caller / pipeline allocation. Every GPU that loads the library holds
// the array (66 560 bytes) until exit. ALL calls on a device share it and it keeps the last call's contents: every word a call
// reads is written by that same call first (k_live_counts writes counts[0 .. blocks), k_scan_counts writes counts[blocks]), so
// nothing has to be cleared — but two calls that overlap on the GPU (different streams, no ordering between them) race on it.
#include <mutex>
#include <string>

#include "../../include/prestogpu.h"

namespace pgpu {
void setLastGlobalError(const std::string& m);  // pipeline.cpp: the text pgpu_last_global_error() returns (thread-local)
}

namespace {
constexpr int64_t kStaticMaxPositions = PGPU_EXPAND_STATIC_MAX_POSITIONS;
constexpr int kStaticScratchWords = (int)(kStaticMaxPositions / pgpu::kSerdeBlockPositions) + 256;  // blocks + total, padded to 256-byte multiple
static_assert((size_t)kStaticScratchWords >= (size_t)(kStaticMaxPositions / pgpu::kSerdeBlockPositions) + 1, "scratch too small for the stated maximum");
static_assert(kStaticScratchWords * 4 % 256 == 0, "scratch size is a multiple of its alignment");

static __device__ __align__(256) uint32_t g_expandScratch[kStaticScratchWords];

constexpr int kStaticMaxDevices = 64;
std::mutex g_scratchMu;
uint32_t* g_scratchAddr[kStaticMaxDevices];  // device address of g_expandScratch on device d (the symbol itself is a host-side stand-in)
}  // namespace

extern "C" pgpu_status pgpu_expand_nulls_static(int32_t device, void* stream, const uint8_t* null_bits, int64_t positions, const void* compact_values,
                                                int32_t width, void* values_out, uint8_t* nulls_out) {
    auto fail = [](pgpu_status st, const std::string& m) {
        pgpu::setLastGlobalError("pgpu_expand_nulls_static: " + m);
        return st;
    };
    if (device < 0 || device >= kStaticMaxDevices) return fail(PGPU_INVALID_ARGUMENT, "device index out of range");
    if (positions < 0) return fail(PGPU_INVALID_ARGUMENT, "negative position count");
    if (positions > kStaticMaxPositions)
        return fail(PGPU_INVALID_ARGUMENT, "position count " + std::to_string(positions) + " exceeds the static scratch's maximum of " +
                                               std::to_string(kStaticMaxPositions) + " (use pgpu_add_serialized_page, which sizes its own scratch)");
    if (width != 1 && width != 4 && width != 8 && width != 16) return fail(PGPU_INVALID_ARGUMENT, "width must be 1, 4, 8 or 16 bytes");
    if (positions == 0) return PGPU_OK;
    if (!null_bits || !compact_values || !values_out || !nulls_out) return fail(PGPU_INVALID_ARGUMENT, "null argument");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(PGPU_CUDA_ERROR, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    uint32_t* scratch = nullptr;
    {
        // once per device, with that device current: the device address of the array
        std::lock_guard<std::mutex> lk(g_scratchMu);
        if (!g_scratchAddr[device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_expandScratch);
            if (e != cudaSuccess || !p) return fail(PGPU_CUDA_ERROR, std::string("cudaGetSymbolAddress: ") + cudaGetErrorString(e));
            g_scratchAddr[device] = (uint32_t*)p;
        }
        scratch = g_scratchAddr[device];
    }
    e = pgpu::expandNulls((cudaStream_t)stream, null_bits, positions, compact_values, values_out, nulls_out, width, scratch);
    if (e != cudaSuccess) return fail(PGPU_CUDA_ERROR, std::string("launch failed: ") + cudaGetErrorString(e));
    return PGPU_OK;
}
