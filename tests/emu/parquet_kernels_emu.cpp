// TEST INFRASTRUCTURE (never linked into the product): compiles the CUDA kernels of presto_b200/csrc/parquet.cu as plain C++ and
// runs them thread by thread on the CPU — the same source the GPU executes, with blockIdx / threadIdx / gridDim provided here and
// the <<<>>> launch wrappers (excluded by PGPU_HOST_EMULATION) restated as loops over the grid. None of these kernels uses shared
// memory or barriers, so a sequential walk over the threads is a faithful execution. tests/test_parquet_kernels_emu.py checks the
// results against the oracle; the GPU suite (tests/test_gpu_parquet.py) checks the real thing.
#define PGPU_HOST_EMULATION 1
#include <cuda_runtime.h>  // vector types (ulonglong2, dim3); host side only
#include <stdint.h>

#undef __global__
#undef __device__
#undef __launch_bounds__
#define __global__
#define __device__
#define __launch_bounds__(...)

static thread_local uint3 threadIdx, blockIdx;
static thread_local dim3 gridDim, blockDim;
static inline unsigned atomicOr(unsigned* p, unsigned v) {
    const unsigned old = *p;
    *p |= v;
    return old;
}

#include "../../presto_b200/csrc/parquet.cu"

namespace {

template <typename F>
void emuLaunch(int grid, int block, F body) {
    gridDim = dim3((unsigned)grid);
    blockDim = dim3((unsigned)block);
    for (int b = 0; b < grid; b++)
        for (int t = 0; t < block; t++) {
            blockIdx.x = (unsigned)b;
            threadIdx.x = (unsigned)t;
            body();
        }
}

}  // namespace

using namespace pgpu;

extern "C" {

// the launch geometry and dispatch of pqLevelsToNullBits / pqExpandRuns / pqConvertPlain / pqNullIds (parquet.cu), as loops
int emu_levels_to_null_bits(const uint8_t* chunk, size_t chunkLen, const PqRun* runs, int64_t nruns, int64_t n, uint8_t* bits) {
    if (n <= 0) return 0;
    emuLaunch(gridFor((n + 7) / 8), kThreads, [&] { k_pq_null_bits(chunk, chunkLen, runs, nruns, n, bits); });
    return 0;
}

int emu_expand_runs(const uint8_t* chunk, size_t chunkLen, const PqRun* runs, int64_t nruns, int64_t first, int64_t count, int outWidth, const void* dict,
                    uint32_t limit, void* out, uint32_t* err) {
    if (count <= 0) return 0;
    const int g = gridFor((count + kPer - 1) / kPer);
    if (dict) {
        if (outWidth == 16) emuLaunch(g, kThreads, [&] { k_pq_expand<ulonglong2, true>(chunk, chunkLen, runs, nruns, first, count, (const ulonglong2*)dict, limit, (ulonglong2*)out, err); });
        else if (outWidth == 8) emuLaunch(g, kThreads, [&] { k_pq_expand<uint64_t, true>(chunk, chunkLen, runs, nruns, first, count, (const uint64_t*)dict, limit, (uint64_t*)out, err); });
        else if (outWidth == 4) emuLaunch(g, kThreads, [&] { k_pq_expand<uint32_t, true>(chunk, chunkLen, runs, nruns, first, count, (const uint32_t*)dict, limit, (uint32_t*)out, err); });
        else return 1;
    } else {
        if (outWidth == 4) emuLaunch(g, kThreads, [&] { k_pq_expand<uint32_t, false>(chunk, chunkLen, runs, nruns, first, count, nullptr, limit, (uint32_t*)out, err); });
        else if (outWidth == 1) emuLaunch(g, kThreads, [&] { k_pq_expand<uint8_t, false>(chunk, chunkLen, runs, nruns, first, count, nullptr, limit, (uint8_t*)out, err); });
        else return 1;
    }
    return 0;
}

int emu_convert_plain(const uint8_t* src, int inWidth, int kind, int64_t count, void* out, int outWidth) {
    if (count <= 0) return 0;
    const int g = gridFor(count);
    if (kind == 0 && inWidth == 4 && outWidth == 8) emuLaunch(g, kThreads, [&] { k_pq_convert<0, 8>(src, inWidth, count, out); });
    else if (kind == 1 && inWidth >= 1 && inWidth <= 8 && outWidth == 8) emuLaunch(g, kThreads, [&] { k_pq_convert<1, 8>(src, inWidth, count, out); });
    else if (kind == 1 && inWidth >= 1 && inWidth <= 16 && outWidth == 16) emuLaunch(g, kThreads, [&] { k_pq_convert<1, 16>(src, inWidth, count, out); });
    else return 1;
    return 0;
}

int emu_null_ids(int32_t* ids, const uint8_t* nulls, int64_t n, int32_t nullId) {
    if (n <= 0) return 0;
    emuLaunch(gridFor(n), kThreads, [&] { k_pq_null_ids(ids, nulls, n, nullId); });
    return 0;
}

}  // extern "C"
