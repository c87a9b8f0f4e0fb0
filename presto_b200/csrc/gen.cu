// gen.cu — ahead-of-time sm_100a kernels that do not depend on a plan: synthetic lineitem columns generated
// straight into HBM (bench / parity inputs) and the L2 flush used between timed iterations.
#include <cuda_runtime.h>

#include <thread>
#include <vector>

#include "../../include/prestogpu_bench.h"
#include "hostcopy.h"
#include "lineitem_gen.h"
#include "tpch_dbgen.cuh"

namespace {

__global__ void __launch_bounds__(256) k_gen_lineitem(int column, unsigned long long seed, long long first, long long rows, long long total, void* out) {
    long long n = rows + (column == PGPU_LI_IOTA_OFFSETS ? 1 : 0);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        ligen::store(column, seed, first + i, first, total, out);
}

__global__ void __launch_bounds__(256) k_flush(uint4* p, size_t n16, unsigned v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) p[i] = make_uint4(v, v, v, v);
}

// ---- TPC-H rows (tpch_dbgen.cuh): one thread per order --------------------------------------------------------------------
constexpr int kGenThreads = 256;

__host__ __device__ inline bool tpchColumnOk(int column, int format) {
    if (format != PGPU_TPCH_NATIVE && format != PGPU_TPCH_DECIMAL_12_2) return false;
    if (format == PGPU_TPCH_DECIMAL_12_2) return column >= PGPU_LI_QUANTITY && column <= PGPU_LI_TAX;
    return column >= PGPU_LI_ORDERKEY && column <= PGPU_LI_LINESTATUS_ID && column != PGPU_LI_IOTA_OFFSETS;
}

__host__ __device__ inline void tpchStore(int column, int format, const dbgen::Line& l, long long orderIndex, long long row, void* out) {
    const bool dec = format == PGPU_TPCH_DECIMAL_12_2;
    switch (column) {
        case PGPU_LI_ORDERKEY: ((long long*)out)[row] = dbgen::sparseOrderKey(orderIndex); break;
        case PGPU_LI_QUANTITY:
            if (dec) ((long long*)out)[row] = (long long)l.quantity * 100; else ((double*)out)[row] = (double)l.quantity;
            break;
        case PGPU_LI_EXTENDEDPRICE:
            if (dec) ((long long*)out)[row] = l.extendedPrice; else ((double*)out)[row] = (double)l.extendedPrice / 100.0;
            break;
        case PGPU_LI_DISCOUNT:
            if (dec) ((long long*)out)[row] = l.discount; else ((double*)out)[row] = (double)l.discount / 100.0;
            break;
        case PGPU_LI_TAX:
            if (dec) ((long long*)out)[row] = l.tax; else ((double*)out)[row] = (double)l.tax / 100.0;
            break;
        case PGPU_LI_SHIPDATE: ((int*)out)[row] = l.shipDate; break;
        case PGPU_LI_RETURNFLAG: ((unsigned char*)out)[row] = (unsigned char)l.returnFlag; break;
        case PGPU_LI_LINESTATUS: ((unsigned char*)out)[row] = (unsigned char)l.lineStatus; break;
        case PGPU_LI_DISCOUNT_ID: ((int*)out)[row] = l.discount; break;
        case PGPU_LI_TAX_ID: ((int*)out)[row] = l.tax; break;
        case PGPU_LI_QUANTITY_ID: ((int*)out)[row] = l.quantity - 1; break;
        case PGPU_LI_RETURNFLAG_ID: ((int*)out)[row] = l.returnFlag == 'A' ? 0 : l.returnFlag == 'N' ? 1 : 2; break;
        case PGPU_LI_LINESTATUS_ID: ((int*)out)[row] = l.lineStatus == 'F' ? 0 : 1; break;
        default: break;
    }
}

// rows of each block of kGenThreads orders
__global__ void __launch_bounds__(kGenThreads) k_tpch_block_rows(long long first, long long n, unsigned* blockRows) {
    __shared__ unsigned s_total;
    if (threadIdx.x == 0) s_total = 0;
    __syncthreads();
    const long long i = (long long)blockIdx.x * kGenThreads + threadIdx.x;
    unsigned c = i < n ? (unsigned)dbgen::lineCountAt(first + i) : 0u;
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0) atomicAdd(&s_total, c);
    __syncthreads();
    if (threadIdx.x == 0) blockRows[blockIdx.x] = s_total;
}

// base[i] = sum of rows[0 .. i), base[m] = total; one block of 1024 threads
__global__ void __launch_bounds__(1024) k_tpch_scan(const unsigned* rows, unsigned long long* base, long long m) {
    __shared__ unsigned long long s_warp[32];
    __shared__ unsigned long long s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long long i0 = 0; i0 < m; i0 += 1024) {
        const long long i = i0 + threadIdx.x;
        const unsigned long long v = i < m ? rows[i] : 0ull;
        unsigned long long incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const unsigned long long w = s_warp[lane];
            unsigned long long wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned long long t = __shfl_up_sync(0xffffffffu, wi, d);
                if (lane >= d) wi += t;
            }
            s_warp[lane] = wi - w;
        }
        __syncthreads();
        const unsigned long long carry = s_carry;
        if (i < m) base[i] = carry + s_warp[warp] + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + s_warp[warp] + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) base[m] = s_carry;
}

__global__ void __launch_bounds__(kGenThreads) k_tpch_lineitem(double sf, long long first, long long n, int column, int format,
                                                                 const unsigned long long* blockBase, long long capacity, void* out) {
    __shared__ unsigned s_warp[kGenThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long i = (long long)blockIdx.x * kGenThreads + threadIdx.x;
    dbgen::Order o;
    o.lines = 0;
    if (i < n) o = dbgen::orderAt(first + i);
    unsigned incl = (unsigned)o.lines;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    unsigned before = 0;
    for (int w = 0; w < warp; w++) before += s_warp[w];
    if (i >= n) return;
    long long row = (long long)blockBase[blockIdx.x] + before + incl - (unsigned)o.lines;
    dbgen::LineCursor cur(o, sf);
    for (int ln = 0; ln < o.lines; ln++, row++) {
        const dbgen::Line l = cur.next();
        if (row < capacity) tpchStore(column, format, l, first + i + 1, row, out);
    }
}

unsigned hostThreads(long long work) {
    unsigned nt = std::max(1u, std::min(std::thread::hardware_concurrency(), 64u));
    return work < (1 << 14) ? 1u : nt;
}

// rows of orders [first, first + n): the line-count stream is stepped, not re-derived per order
long long tpchRowsHost(long long first, long long n) {
    const unsigned nt = hostThreads(n);
    std::vector<long long> part(nt, 0);
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++)
        th.emplace_back([=, &part]() {
            const long long a = n * t / nt, b = n * (t + 1) / nt;
            uint32_t s = dbgen::mulmod(dbgen::kSeedLineCount, dbgen::powmod((uint64_t)(first + a)));
            long long c = 0;
            for (long long i = a; i < b; i++) c += dbgen::draw(s, 1, dbgen::kMaxLines);
            part[t] = c;
        });
    for (auto& x : th) x.join();
    long long total = 0;
    for (long long c : part) total += c;
    return total;
}

}  // namespace

extern "C" {

int64_t pgpu_tpch_order_count(double sf) { return sf > 0 ? dbgen::orderCount(sf) : 0; }

pgpu_status pgpu_tpch_lineitem_rows(double sf, int64_t first_order, int64_t n_orders, int64_t* rows) {
    if (!rows || first_order < 0 || n_orders < 0 || first_order + n_orders > dbgen::orderCount(sf)) return PGPU_INVALID_ARGUMENT;
    *rows = tpchRowsHost(first_order, n_orders);
    return PGPU_OK;
}

pgpu_status pgpu_tpch_lineitem_device(int32_t device, double sf, int64_t first_order, int64_t n_orders, int32_t column, int32_t format,
                                      int64_t capacity_rows, void* dev_out, int64_t* rows) {
    if (!tpchColumnOk(column, format) || first_order < 0 || n_orders < 0 || first_order + n_orders > dbgen::orderCount(sf) || capacity_rows < 0 ||
        (!dev_out && capacity_rows))
        return PGPU_INVALID_ARGUMENT;
    if (rows) *rows = 0;
    if (n_orders == 0) return PGPU_OK;
    if (cudaSetDevice(device) != cudaSuccess) return PGPU_CUDA_ERROR;
    const long long blocks = (n_orders + kGenThreads - 1) / kGenThreads;
    unsigned* blockRows = nullptr;
    unsigned long long* blockBase = nullptr;
    if (cudaMalloc(&blockRows, (size_t)blocks * 4) != cudaSuccess) return PGPU_INSUFFICIENT_RESOURCES;
    if (cudaMalloc(&blockBase, (size_t)(blocks + 1) * 8) != cudaSuccess) {
        cudaFree(blockRows);
        return PGPU_INSUFFICIENT_RESOURCES;
    }
    k_tpch_block_rows<<<(unsigned)blocks, kGenThreads>>>(first_order, n_orders, blockRows);
    k_tpch_scan<<<1, 1024>>>(blockRows, blockBase, blocks);
    unsigned long long total = 0;
    cudaError_t e = cudaMemcpy(&total, blockBase + blocks, 8, cudaMemcpyDeviceToHost);
    pgpu_status st = PGPU_OK;
    if (e != cudaSuccess) st = PGPU_CUDA_ERROR;
    else if ((long long)total > capacity_rows) st = PGPU_INVALID_ARGUMENT;  // the caller sizes buffers with pgpu_tpch_lineitem_rows
    if (!st) {
        k_tpch_lineitem<<<(unsigned)blocks, kGenThreads>>>(sf, first_order, n_orders, column, format, blockBase, capacity_rows, dev_out);
        if (cudaGetLastError() != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) st = PGPU_CUDA_ERROR;
    }
    cudaFree(blockRows);
    cudaFree(blockBase);
    if (rows) *rows = (int64_t)total;
    return st;
}

pgpu_status pgpu_tpch_lineitem_host(double sf, int64_t first_order, int64_t n_orders, int32_t column, int32_t format, int64_t capacity_rows,
                                    void* host_out, int64_t* rows) {
    if (!tpchColumnOk(column, format) || first_order < 0 || n_orders < 0 || first_order + n_orders > dbgen::orderCount(sf) || capacity_rows < 0 ||
        (!host_out && capacity_rows))
        return PGPU_INVALID_ARGUMENT;
    const unsigned nt = hostThreads(n_orders);
    std::vector<long long> start(nt + 1, 0);
    for (unsigned t = 0; t < nt; t++) start[t + 1] = start[t] + tpchRowsHost(first_order + n_orders * t / nt, n_orders * (t + 1) / nt - n_orders * t / nt);
    if (rows) *rows = start[nt];
    if (start[nt] > capacity_rows) return PGPU_INVALID_ARGUMENT;
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++)
        th.emplace_back([=, &start]() {
            long long row = start[t];
            for (long long i = n_orders * t / nt; i < n_orders * (t + 1) / nt; i++) {
                const dbgen::Order o = dbgen::orderAt(first_order + i);
                dbgen::LineCursor cur(o, sf);
                for (int ln = 0; ln < o.lines; ln++, row++) tpchStore(column, format, cur.next(), first_order + i + 1, row, host_out);
            }
        });
    for (auto& x : th) x.join();
    return PGPU_OK;
}

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

pgpu_status pgpu_memcpy_h2d(int32_t device, void* dev_dst, const void* host_src, size_t bytes) {
    if (cudaSetDevice(device) != cudaSuccess) return PGPU_CUDA_ERROR;
    return cudaMemcpy(dev_dst, host_src, bytes, cudaMemcpyHostToDevice) == cudaSuccess ? PGPU_OK : PGPU_CUDA_ERROR;
}

pgpu_status pgpu_memcpy_d2h(int32_t device, void* host_dst, const void* dev_src, size_t bytes) {
    if (cudaSetDevice(device) != cudaSuccess) return PGPU_CUDA_ERROR;
    return cudaMemcpy(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost) == cudaSuccess ? PGPU_OK : PGPU_CUDA_ERROR;
}

// the library's staging copy alone (pageable -> pinned, worker pool, non-temporal stores): bench.py measures with it what the host
// memory system gives the ingest path when no GPU is involved
pgpu_status pgpu_bench_host_copy(void* dst, const void* src, size_t bytes) {
    if (!dst || !src) return PGPU_INVALID_ARGUMENT;
    std::vector<pgpu::CopyTask> t;
    t.push_back(pgpu::CopyTask{pgpu::CopyTask::BYTES, dst, src, bytes, 0});
    pgpu::hostCopy(t);
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
