// tableutil.cu — see tableutil.h.
#include "tableutil.h"

namespace pgpu {
namespace {

template <typename T>
__global__ void k_key_clustering(const T* __restrict__ keys, long long n, unsigned long long* out) {
    unsigned changes = 0, near[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (long long i = 1 + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const long long a = (long long)keys[i - 1], b = (long long)keys[i];
        if (a != b) {
            changes++;
#pragma unroll
            for (int d = 0; d < 8; d++) near[d] += ((a + d) >> 3) == ((b + d) >> 3) ? 1u : 0u;
        }
    }
    changes = __reduce_add_sync(0xffffffffu, changes);
#pragma unroll
    for (int d = 0; d < 8; d++) near[d] = __reduce_add_sync(0xffffffffu, near[d]);
    if ((threadIdx.x & 31) == 0) {
        if (changes) atomicAdd(out, (unsigned long long)changes);
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

}  // namespace

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
