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

}  // namespace

cudaError_t keyClustering(cudaStream_t stream, const void* keys, int width, int64_t n, unsigned long long* out) {
    if (n < 2) return cudaSuccess;
    const int block = 256, grid = (int)((n + block - 1) / block < 64 ? (n + block - 1) / block : 64);
    if (width == 4) k_key_clustering<int><<<grid, block, 0, stream>>>((const int*)keys, (long long)n, out);
    else k_key_clustering<long long><<<grid, block, 0, stream>>>((const long long*)keys, (long long)n, out);
    return cudaGetLastError();
}

}  // namespace pgpu
