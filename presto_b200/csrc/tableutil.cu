// tableutil.cu — see tableutil.h.
#include "tableutil.h"

namespace pgpu {
namespace {

template <typename T>
__global__ void k_key_clustering(const T* __restrict__ keys, long long n, unsigned long long* out) {
    unsigned changes = 0, descents = 0, near[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (long long i = 1 + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const long long a = (long long)keys[i - 1], b = (long long)keys[i];
        if (b < a) descents++;  // the order the streaming-runs path expects of its first grouping field (P::key_order: signed)
        if (a != b) {
            changes++;
#pragma unroll
            for (int d = 0; d < 8; d++) near[d] += ((a + d) >> 3) == ((b + d) >> 3) ? 1u : 0u;
        }
    }
    changes = __reduce_add_sync(0xffffffffu, changes);
    descents = __reduce_add_sync(0xffffffffu, descents);
#pragma unroll
    for (int d = 0; d < 8; d++) near[d] = __reduce_add_sync(0xffffffffu, near[d]);
    if ((threadIdx.x & 31) == 0) {
        if (changes) atomicAdd(out, (unsigned long long)changes);
        if (descents) atomicAdd(out + 9, (unsigned long long)descents);
#pragma unroll
        for (int d = 0; d < 8; d++)
            if (near[d]) atomicAdd(out + 1 + d, (unsigned long long)near[d]);
    }
}

// one block of 1024 threads, one partition per thread
__global__ void k_partition_scan(const unsigned* __restrict__ hist, unsigned long long* cursor, unsigned long long* start, int K) {
    __shared__ unsigned long long warpSum[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned long long v = (int)threadIdx.x < K ? hist[threadIdx.x] : 0ull;
    unsigned long long x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
    }
    if (lane == 31) warpSum[warp] = x;
    __syncthreads();
    if (warp == 0) {
        unsigned long long w = warpSum[lane], wx = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xffffffffu, wx, d);
            if (lane >= d) wx += y;
        }
        warpSum[lane] = wx - w;  // exclusive
    }
    __syncthreads();
    const unsigned long long excl = warpSum[warp] + x - v;
    if ((int)threadIdx.x < K) {
        cursor[threadIdx.x] = excl;
        start[threadIdx.x] = excl;
        if ((int)threadIdx.x == K - 1) start[K] = excl + v;
    }
}

// exclusive scan of `n` 32-bit counts into 64-bit offsets, two launches: every block sums its segment of `per` counts, then every
// block adds up the sums of the blocks before it and scans its own segment in strips of blockDim.x
constexpr int kScanThreads = 256;
__device__ unsigned long long blockSum(unsigned long long v, unsigned long long* warpSum) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
    __syncthreads();  // warpSum may still be read from the call before
    if (lane == 0) warpSum[warp] = v;
    __syncthreads();
    unsigned long long t = 0;
    for (int w = 0; w < kScanThreads / 32; w++) t += warpSum[w];
    return t;
}
__global__ void __launch_bounds__(kScanThreads) k_runs_sums(const unsigned* counts, long long n, long long per, unsigned long long* sums) {
    __shared__ unsigned long long warpSum[kScanThreads / 32];
    const long long b = (long long)blockIdx.x * per, e = b + per < n ? b + per : n;
    unsigned long long v = 0;
    for (long long i = b + threadIdx.x; i < e; i += kScanThreads) v += counts[i];
    v = blockSum(v, warpSum);
    if (threadIdx.x == 0) sums[blockIdx.x] = v;
}
__global__ void __launch_bounds__(kScanThreads) k_runs_offsets(const unsigned* counts, long long n, long long per, const unsigned long long* sums,
                                                               unsigned long long* offsets, unsigned long long* total) {
    __shared__ unsigned long long warpSum[kScanThreads / 32];
    unsigned long long v = 0;
    for (int j = threadIdx.x; j < (int)blockIdx.x; j += kScanThreads) v += sums[j];
    unsigned long long base = blockSum(v, warpSum);
    const long long b = (long long)blockIdx.x * per, e = b + per < n ? b + per : n;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long long s0 = b; s0 < e; s0 += kScanThreads) {
        const long long i = s0 + threadIdx.x;
        const unsigned long long c = i < e ? counts[i] : 0ull;
        unsigned long long x = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        __syncthreads();
        if (lane == 31) warpSum[warp] = x;
        __syncthreads();
        unsigned long long before = 0, strip = 0;
        for (int w = 0; w < kScanThreads / 32; w++) {
            const unsigned long long t = warpSum[w];
            before += w < warp ? t : 0ull;
            strip += t;
        }
        if (i < e) offsets[i] = base + before + x - c;
        base += strip;
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) *total = base;
}

}  // namespace

cudaError_t runsScan(cudaStream_t stream, const unsigned* counts, long long n, unsigned long long* offsets, unsigned long long* sums, int nblocks,
                     unsigned long long* total) {
    if (n <= 0) return cudaMemsetAsync(total, 0, 8, stream);
    if (nblocks > 1024) nblocks = 1024;
    long long per = (n + nblocks - 1) / nblocks;
    per = (per + kScanThreads - 1) / kScanThreads * kScanThreads;
    nblocks = (int)((n + per - 1) / per);
    k_runs_sums<<<nblocks, kScanThreads, 0, stream>>>(counts, n, per, sums);
    k_runs_offsets<<<nblocks, kScanThreads, 0, stream>>>(counts, n, per, sums, offsets, total);
    return cudaGetLastError();
}

cudaError_t partitionScan(cudaStream_t stream, const unsigned* hist, unsigned long long* cursor, unsigned long long* start, int K) {
    if (K < 1 || K > 1024) return cudaErrorInvalidValue;
    k_partition_scan<<<1, 1024, 0, stream>>>(hist, cursor, start, K);
    return cudaGetLastError();
}

cudaError_t keyClustering(cudaStream_t stream, const void* keys, int width, int64_t n, unsigned long long* out) {
    if (n < 2) return cudaSuccess;
    const int block = 256, grid = (int)((n + block - 1) / block < 64 ? (n + block - 1) / block : 64);
    if (width == 4) k_key_clustering<int><<<grid, block, 0, stream>>>((const int*)keys, (long long)n, out);
    else k_key_clustering<long long><<<grid, block, 0, stream>>>((const long long*)keys, (long long)n, out);
    return cudaGetLastError();
}

}  // namespace pgpu
