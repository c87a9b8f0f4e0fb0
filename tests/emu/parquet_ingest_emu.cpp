// TEST INFRASTRUCTURE (never linked into the product): the whole device side of the Parquet ingest on the CPU. Compiles the
// product's orchestration (presto_b200/csrc/parquet_ingest.cpp: arena layout, copies, launch sequence) and its kernels
// (parquet.cu) as plain C++: "device memory" is host memory, cudaMemcpyAsync / cudaMemsetAsync are memcpy / memset, kernel
// launches are loops over the grid. The one piece restated rather than compiled is serde.cu's expandNulls (warp shuffles and
// shared memory; it has run on the GPU since the wire-format ingest) — from its contract in serde.h.
// tests/test_parquet_ingest_emu.py compares the resulting Blocks with the oracle.
#define PGPU_HOST_EMULATION 1
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

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
// the CUDA runtime calls of parquet_ingest.cpp, on host memory
#define cudaMemcpyAsync(dst, src, n, kind, stream) (memcpy((dst), (src), (n)), cudaSuccess)
#define cudaMemsetAsync(dst, v, n, stream) (memset((dst), (v), (n)), cudaSuccess)
#define cudaGetErrorString(e) "emulated"

#include "../../presto_b200/csrc/parquet.cu"

namespace pgpu {
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

// the launch wrappers of parquet.cu (excluded there by PGPU_HOST_EMULATION): same geometry and dispatch, loops for launches
cudaError_t pqLevelsToNullBits(cudaStream_t, const uint8_t* chunk, size_t chunkLen, const PqRun* runs, int64_t nruns, int64_t n, uint8_t* bits) {
    if (n <= 0) return cudaSuccess;
    emuLaunch(gridFor((n + 7) / 8), kThreads, [&] { k_pq_null_bits(chunk, chunkLen, runs, nruns, n, bits); });
    return cudaSuccess;
}
cudaError_t pqExpandRuns(cudaStream_t, const uint8_t* chunk, size_t chunkLen, const PqRun* runs, int64_t nruns, int64_t first, int64_t count, int outWidth, const void* dict,
                         uint32_t limit, void* out, uint32_t* err) {
    if (count <= 0) return cudaSuccess;
    const int g = gridFor((count + kPer - 1) / kPer);
    if (dict) {
        if (outWidth == 16) emuLaunch(g, kThreads, [&] { k_pq_expand<ulonglong2, true>(chunk, chunkLen, runs, nruns, first, count, (const ulonglong2*)dict, limit, (ulonglong2*)out, err); });
        else if (outWidth == 8) emuLaunch(g, kThreads, [&] { k_pq_expand<uint64_t, true>(chunk, chunkLen, runs, nruns, first, count, (const uint64_t*)dict, limit, (uint64_t*)out, err); });
        else if (outWidth == 4) emuLaunch(g, kThreads, [&] { k_pq_expand<uint32_t, true>(chunk, chunkLen, runs, nruns, first, count, (const uint32_t*)dict, limit, (uint32_t*)out, err); });
        else return cudaErrorInvalidValue;
    } else {
        if (outWidth == 4) emuLaunch(g, kThreads, [&] { k_pq_expand<uint32_t, false>(chunk, chunkLen, runs, nruns, first, count, nullptr, limit, (uint32_t*)out, err); });
        else if (outWidth == 1) emuLaunch(g, kThreads, [&] { k_pq_expand<uint8_t, false>(chunk, chunkLen, runs, nruns, first, count, nullptr, limit, (uint8_t*)out, err); });
        else return cudaErrorInvalidValue;
    }
    return cudaSuccess;
}
cudaError_t pqConvertPlain(cudaStream_t, const uint8_t* src, int inWidth, int kind, int64_t count, void* out, int outWidth) {
    if (count <= 0) return cudaSuccess;
    const int g = gridFor(count);
    if (kind == 0 && inWidth == 4 && outWidth == 8) emuLaunch(g, kThreads, [&] { k_pq_convert<0, 8>(src, inWidth, count, out); });
    else if (kind == 1 && inWidth >= 1 && inWidth <= 8 && outWidth == 8) emuLaunch(g, kThreads, [&] { k_pq_convert<1, 8>(src, inWidth, count, out); });
    else if (kind == 1 && inWidth >= 1 && inWidth <= 16 && outWidth == 16) emuLaunch(g, kThreads, [&] { k_pq_convert<1, 16>(src, inWidth, count, out); });
    else return cudaErrorInvalidValue;
    return cudaSuccess;
}
cudaError_t pqNullIds(cudaStream_t, int32_t* ids, const uint8_t* nulls, int64_t n, int32_t nullId) {
    if (n <= 0) return cudaSuccess;
    emuLaunch(gridFor(n), kThreads, [&] { k_pq_null_ids(ids, nulls, n, nullId); });
    return cudaSuccess;
}
// serde.h: bits (most significant bit first, 1 = NULL) -> nulls[n]; compact[non-null] -> values[n], zero at NULL positions
cudaError_t expandNulls(cudaStream_t, const uint8_t* bits, int64_t n, const void* compact, void* values, uint8_t* nulls, int width, uint32_t*) {
    int64_t src = 0;
    for (int64_t i = 0; i < n; i++) {
        const bool isNull = (bits[i >> 3] >> (7 - (int)(i & 7))) & 1;
        nulls[i] = isNull ? 1 : 0;
        if (isNull) memset((char*)values + i * width, 0, (size_t)width);
        else memcpy((char*)values + i * width, (const char*)compact + (src++) * width, (size_t)width);
    }
    return cudaSuccess;
}
}  // namespace pgpu

#include "../../presto_b200/csrc/parquet_ingest.cpp"

using namespace pgpu;

namespace {
struct Session {
    std::unique_ptr<Plan> plan;
    PqDecoded dec;
    char* arena = nullptr;
    uint32_t err = 0;
    std::string why;
};
}  // namespace

extern "C" {

// -> a session holding the decoded columns, or nullptr (then *status / message say why)
void* emu_decode_row_group(const char* planText, const pgpu_parquet_chunk* chunks, int32_t nchunks, int64_t rows, int pageThreads, int32_t* status, char* message, size_t messageLen) {
    Session* s = new Session();
    *status = PGPU_OK;
    try {
        s->plan = parsePlan(planText);
        *status = pqDecodeRowGroup(*s->plan, chunks, nchunks, rows, nullptr, &s->err, pageThreads, [&](size_t need) -> char* {
            s->arena = (char*)malloc(need);   // exactly `need` bytes: tools like valgrind / ASan see any access past the layout
            return s->arena;
        }, s->dec, s->why);
    } catch (const PlanError& e) {
        *status = e.status;
        s->why = e.what();
    }
    if (*status != PGPU_OK) {
        if (message && messageLen) {
            strncpy(message, s->why.c_str(), messageLen - 1);
            message[messageLen - 1] = 0;
        }
        free(s->arena);
        delete s;
        return nullptr;
    }
    return s;
}
const pgpu_column* emu_column(void* session, int32_t c) { return &((Session*)session)->dec.cols[(size_t)c]; }
uint32_t emu_error_bits(void* session) { return ((Session*)session)->err; }
int32_t emu_launches(void* session) { return ((Session*)session)->dec.launches; }
void emu_free(void* session) {
    Session* s = (Session*)session;
    free(s->arena);
    delete s;
}

}  // extern "C"
