// parquet.cu — see parquet.h. Ahead-of-time sm_100a kernels.
#include "parquet.h"

namespace pgpu {
namespace {

constexpr int kThreads = 256;
constexpr uint32_t kPacked = 0x80000000u;

// index of the run that holds value `idx` (runs sorted by start, gap-free from runs[0].start)
__device__ inline int64_t findRun(const PqRun* runs, int64_t nruns, int64_t idx) {
    int64_t lo = 0, hi = nruns;  // first run with start > idx
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (runs[mid].start <= idx) lo = mid + 1;
        else hi = mid;
    }
    return lo - 1;
}

__device__ inline uint32_t codeAt(const uint8_t* chunk, size_t chunkLen, const PqRun& r, int64_t idx) {
    const uint32_t width = r.widthPacked & 0xffu;
    if (!(r.widthPacked & kPacked)) return (uint32_t)r.src;
    if (width == 0) return 0;
    const uint64_t bit = r.src + (uint64_t)(idx - r.start) * width;
    const uint64_t byte = bit >> 3;
    const uint32_t shift = (uint32_t)(bit & 7);
    uint64_t w = 0;
    const int need = (int)((shift + width + 7) >> 3);  // <= 5
    for (int b = 0; b < need; b++)
        if (byte + b < chunkLen) w |= (uint64_t)chunk[byte + b] << (8 * b);
    return (uint32_t)((w >> shift) & (width == 32 ? 0xffffffffull : ((1ull << width) - 1)));
}

__global__ void __launch_bounds__(kThreads) k_pq_null_bits(const uint8_t* chunk, size_t chunkLen, const PqRun* runs, int64_t nruns, int64_t n, uint8_t* bits) {
    const int64_t nbytes = (n + 7) >> 3;
    for (int64_t j = (int64_t)blockIdx.x * kThreads + threadIdx.x; j < nbytes; j += (int64_t)gridDim.x * kThreads) {
        int64_t pos = j << 3;
        int64_t ri = findRun(runs, nruns, pos);
        PqRun r = runs[ri < 0 ? 0 : ri];
        uint32_t out = 0;
        for (int k = 0; k < 8 && pos < n; k++, pos++) {
            while (pos >= r.start + (int64_t)r.count && ri + 1 < nruns) r = runs[++ri];
            const uint32_t level = codeAt(chunk, chunkLen, r, pos);
            if (level == 0) out |= 0x80u >> k;
        }
        bits[j] = (uint8_t)out;
    }
}

template <typename T>
__device__ inline T zeroOf() { return T(0); }
template <>
__device__ inline ulonglong2 zeroOf<ulonglong2>() { return make_ulonglong2(0, 0); }
template <typename T>
__device__ inline T fromCode(uint32_t c) { return (T)c; }
template <>
__device__ inline ulonglong2 fromCode<ulonglong2>(uint32_t c) { return make_ulonglong2(c, 0); }

// Each thread expands kPer consecutive values: one binary search for the first, then a walk along the run table (a run holds
// >= 8 values unless it is a short RLE run), so the search cost is shared and a warp writes kPer * 32 consecutive outputs.
constexpr int kPer = 4;

template <typename OUT, bool GATHER>
__global__ void __launch_bounds__(kThreads) k_pq_expand(const uint8_t* chunk, size_t chunkLen, const PqRun* runs, int64_t nruns, int64_t first, int64_t count,
                                                        const OUT* dict, uint32_t limit, OUT* out, uint32_t* err) {
    bool bad = false;
    const int64_t tiles = (count + kPer - 1) / kPer;
    for (int64_t t = (int64_t)blockIdx.x * kThreads + threadIdx.x; t < tiles; t += (int64_t)gridDim.x * kThreads) {
        int64_t i = t * kPer;
        int64_t ri = findRun(runs, nruns, first + i);
        if (ri < 0) ri = 0;
        PqRun r = runs[ri];
#pragma unroll
        for (int k = 0; k < kPer; k++, i++) {
            if (i >= count) break;
            const int64_t idx = first + i;
            while (idx >= r.start + (int64_t)r.count && ri + 1 < nruns) r = runs[++ri];
            uint32_t code = codeAt(chunk, chunkLen, r, idx);
            if (code >= limit) {
                bad = true;
                code = 0;
            }
            if (GATHER) out[i] = limit ? dict[code] : zeroOf<OUT>();
            else out[i] = fromCode<OUT>(code);
        }
    }
    if (bad) atomicOr(err, kPqErrBadId);
}

template <int KIND, int OUTW>
__global__ void __launch_bounds__(kThreads) k_pq_convert(const uint8_t* src, int inWidth, int64_t count, void* out) {
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < count; i += (int64_t)gridDim.x * kThreads) {
        const uint8_t* s = src + i * inWidth;
        if (KIND == 0) {
            const uint32_t v = (uint32_t)s[0] | (uint32_t)s[1] << 8 | (uint32_t)s[2] << 16 | (uint32_t)s[3] << 24;
            ((int64_t*)out)[i] = (int64_t)(int32_t)v;
        } else {
            // big-endian two's complement, sign-extended to 128 bits
            const bool neg = (s[0] & 0x80) != 0;
            unsigned __int128 v = neg ? ~(unsigned __int128)0 : 0;
            for (int b = 0; b < inWidth; b++) v = (v << 8) | s[b];
            if (OUTW == 8) {
                ((int64_t*)out)[i] = (int64_t)(uint64_t)v;
            } else {
                const unsigned __int128 mag = neg ? (unsigned __int128)0 - v : v;
                uint64_t lo = (uint64_t)mag, hi = (uint64_t)(mag >> 64);
                if (neg) hi |= 0x8000000000000000ull;
                ((ulonglong2*)out)[i] = make_ulonglong2(lo, hi);
            }
        }
    }
}

__global__ void __launch_bounds__(kThreads) k_pq_null_ids(int32_t* ids, const uint8_t* nulls, int64_t n, int32_t nullId) {
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
        if (nulls[i]) ids[i] = nullId;
}

inline int gridFor(int64_t items) {
    const int64_t blocks = (items + kThreads - 1) / kThreads;
    return (int)(blocks < 1 ? 1 : blocks > 148 * 16 ? 148 * 16 : blocks);  // grid-stride loops: at most 16 blocks per SM
}

}  // namespace

#ifndef PGPU_HOST_EMULATION  // tests/emu/parquet_kernels_emu.cpp includes this file to run the kernels above thread by thread on the CPU
cudaError_t pqLevelsToNullBits(cudaStream_t stream, const uint8_t* chunk, size_t chunkLen, const PqRun* runs, int64_t nruns, int64_t n, uint8_t* bits) {
    if (n <= 0) return cudaSuccess;
    k_pq_null_bits<<<gridFor((n + 7) / 8), kThreads, 0, stream>>>(chunk, chunkLen, runs, nruns, n, bits);
    return cudaGetLastError();
}

cudaError_t pqExpandRuns(cudaStream_t stream, const uint8_t* chunk, size_t chunkLen, const PqRun* runs, int64_t nruns, int64_t first, int64_t count, int outWidth,
                         const void* dict, uint32_t limit, void* out, uint32_t* err) {
    if (count <= 0) return cudaSuccess;
    const int g = gridFor((count + kPer - 1) / kPer);
    if (dict) {
        if (outWidth == 16) k_pq_expand<ulonglong2, true><<<g, kThreads, 0, stream>>>(chunk, chunkLen, runs, nruns, first, count, (const ulonglong2*)dict, limit, (ulonglong2*)out, err);
        else if (outWidth == 8) k_pq_expand<uint64_t, true><<<g, kThreads, 0, stream>>>(chunk, chunkLen, runs, nruns, first, count, (const uint64_t*)dict, limit, (uint64_t*)out, err);
        else if (outWidth == 4) k_pq_expand<uint32_t, true><<<g, kThreads, 0, stream>>>(chunk, chunkLen, runs, nruns, first, count, (const uint32_t*)dict, limit, (uint32_t*)out, err);
        else return cudaErrorInvalidValue;
    } else {
        if (outWidth == 4) k_pq_expand<uint32_t, false><<<g, kThreads, 0, stream>>>(chunk, chunkLen, runs, nruns, first, count, nullptr, limit, (uint32_t*)out, err);
        else if (outWidth == 1) k_pq_expand<uint8_t, false><<<g, kThreads, 0, stream>>>(chunk, chunkLen, runs, nruns, first, count, nullptr, limit, (uint8_t*)out, err);
        else return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t pqConvertPlain(cudaStream_t stream, const uint8_t* src, int inWidth, int kind, int64_t count, void* out, int outWidth) {
    if (count <= 0) return cudaSuccess;
    const int g = gridFor(count);
    if (kind == 0 && inWidth == 4 && outWidth == 8) k_pq_convert<0, 8><<<g, kThreads, 0, stream>>>(src, inWidth, count, out);
    else if (kind == 1 && inWidth >= 1 && inWidth <= 8 && outWidth == 8) k_pq_convert<1, 8><<<g, kThreads, 0, stream>>>(src, inWidth, count, out);
    else if (kind == 1 && inWidth >= 1 && inWidth <= 16 && outWidth == 16) k_pq_convert<1, 16><<<g, kThreads, 0, stream>>>(src, inWidth, count, out);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

cudaError_t pqNullIds(cudaStream_t stream, int32_t* ids, const uint8_t* nulls, int64_t n, int32_t nullId) {
    if (n <= 0) return cudaSuccess;
    k_pq_null_ids<<<gridFor(n), kThreads, 0, stream>>>(ids, nulls, n, nullId);
    return cudaGetLastError();
}
#endif  // PGPU_HOST_EMULATION

}  // namespace pgpu
