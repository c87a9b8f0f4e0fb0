// gen.cu — ahead-of-time sm_100a kernels that do not depend on a plan: synthetic lineitem columns generated
// straight into HBM (bench / parity inputs) and the L2 flush used between timed iterations.
#include <cuda_runtime.h>

#include <thread>
#include <vector>

#include "../../include/prestogpu.h"
#include "lineitem_gen.h"

namespace {

__global__ void __launch_bounds__(256) k_gen_lineitem(int column, unsigned long long seed, long long first, long long rows, long long total, void* out) {
    long long n = rows + (column == PGPU_LI_IOTA_OFFSETS ? 1 : 0);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        ligen::store(column, seed, first + i, first, total, out);
}

__global__ void __launch_bounds__(256) k_flush(uint4* p, size_t n16, unsigned v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) p[i] = make_uint4(v, v, v, v);
}

}  // namespace

extern "C" {

pgpu_status pgpu_gen_lineitem_device(int32_t device, int32_t column, uint64_t seed, int64_t first_row, int64_t rows, int64_t total_rows, void* dev_out) {
    if (column < 0 || column > PGPU_LI_ORDERKEY_SHUFFLED || rows < 0 || !dev_out) return PGPU_INVALID_ARGUMENT;
    if (cudaSetDevice(device) != cudaSuccess) return PGPU_CUDA_ERROR;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    k_gen_lineitem<<<sms * 8, 256>>>(column, seed, first_row, rows, total_rows, dev_out);
    if (cudaGetLastError() != cudaSuccess) return PGPU_CUDA_ERROR;
    return cudaDeviceSynchronize() == cudaSuccess ? PGPU_OK : PGPU_CUDA_ERROR;
}

pgpu_status pgpu_gen_lineitem_host(int32_t column, uint64_t seed, int64_t first_row, int64_t rows, int64_t total_rows, void* host_out) {
    if (column < 0 || column > PGPU_LI_ORDERKEY_SHUFFLED || rows < 0 || !host_out) return PGPU_INVALID_ARGUMENT;
    int64_t n = rows + (column == PGPU_LI_IOTA_OFFSETS ? 1 : 0);
    unsigned nt = std::max(1u, std::min(std::thread::hardware_concurrency(), 64u));
    if (n < (1 << 16)) nt = 1;
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++) {
        int64_t a = n * t / nt, b = n * (t + 1) / nt;
        th.emplace_back([=]() {
            for (int64_t i = a; i < b; i++) ligen::store(column, seed, first_row + i, first_row, total_rows, host_out);
        });
    }
    for (auto& t : th) t.join();
    return PGPU_OK;
}

pgpu_status pgpu_flush_l2(int32_t device, void* dev_scratch, size_t bytes) {
    if (!dev_scratch) return PGPU_INVALID_ARGUMENT;
    if (cudaSetDevice(device) != cudaSuccess) return PGPU_CUDA_ERROR;
    static unsigned v = 0;
    k_flush<<<148 * 8, 256>>>((uint4*)dev_scratch, bytes / 16, ++v);
    if (cudaGetLastError() != cudaSuccess) return PGPU_CUDA_ERROR;
    return cudaDeviceSynchronize() == cudaSuccess ? PGPU_OK : PGPU_CUDA_ERROR;
}

}  // extern "C"
