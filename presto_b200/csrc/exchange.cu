// exchange.cu — see exchange.h. Ahead-of-time sm_100a kernels (they do not depend on a plan's expressions).
#include "exchange.h"

namespace pgpu {
namespace {

// spi/type/AbstractLongType.java:126-130
__host__ __device__ inline uint64_t rotl64(uint64_t v, int s) { return (v << s) | (v >> (64 - s)); }
__host__ __device__ inline int64_t hashLong(int64_t value) { return (int64_t)(rotl64((uint64_t)value * 0xC2B2AE3D27D4EB4Full, 31) * 0x9E3779B185EBCA87ull); }
// spi/type/DoubleType.java:163-170 (hash of doubleToLongBits with -0 -> +0; doubleToLongBits canonicalises NaN)
__host__ __device__ inline int64_t hashDoubleBits(int64_t bits) {
    uint64_t u = (uint64_t)bits;
    if ((u << 1) == 0) u = 0;                                                          // +-0 -> +0
    if ((u & 0x7ff0000000000000ull) == 0x7ff0000000000000ull && (u & 0x000fffffffffffffull)) u = 0x7ff8000000000000ull;  // NaN
    return hashLong((int64_t)u);
}
__host__ __device__ inline int32_t partitionOfHash(int64_t rawHash, int nparts) {
    return (int32_t)((uint64_t)(rawHash & 0x7fffffffffffffffll) % (uint64_t)nparts);
}

__device__ inline int32_t partitionOfRow(const PartKeys& keys, int64_t row) {
    int64_t h = 0;  // HashGenerationOptimizer.INITIAL_HASH_VALUE
    for (int k = 0; k < keys.nkeys; k++) {
        const PartKey& pk = keys.k[k];
        int64_t hc = 0;  // TypeUtils.NULL_HASH_CODE
        if (!(pk.nulls && pk.nulls[row])) {
            if (pk.kind == PART_KEY_I64) hc = hashLong(((const int64_t*)pk.values)[row]);
            else if (pk.kind == PART_KEY_I32) hc = hashLong((int64_t)((const int32_t*)pk.values)[row]);
            else hc = hashDoubleBits(((const int64_t*)pk.values)[row]);
        }
        h = (int64_t)(31ull * (uint64_t)h + (uint64_t)hc);
    }
    return partitionOfHash(h, keys.nparts);
}

__global__ void __launch_bounds__(256) k_partition_count(PartKeys keys, int64_t rows, uint8_t* pid, unsigned long long* counts) {
    __shared__ unsigned int hist[kMaxParts];
    for (int i = threadIdx.x; i < kMaxParts; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
        int32_t p = partitionOfRow(keys, r);
        pid[r] = (uint8_t)p;
        atomicAdd(&hist[p], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < keys.nparts; i += blockDim.x)
        if (hist[i]) atomicAdd(&counts[i], (unsigned long long)hist[i]);
}

// Every block reserves one contiguous range per partition for its rows (one global atomic per block and partition),
// then ranks its rows inside the range through shared-memory cursors.
__global__ void __launch_bounds__(256) k_partition_assign(int nparts, int64_t rows, const uint8_t* pid, unsigned long long* cursors, int64_t* dst) {
    __shared__ unsigned int hist[kMaxParts];
    __shared__ unsigned long long base[kMaxParts];
    const int64_t per = (rows + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = r0 + per < rows ? r0 + per : rows;
    for (int i = threadIdx.x; i < kMaxParts; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int64_t r = r0 + threadIdx.x; r < r1; r += blockDim.x) atomicAdd(&hist[pid[r]], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) {
        base[i] = hist[i] ? atomicAdd(&cursors[i], (unsigned long long)hist[i]) : 0ull;
        hist[i] = 0;
    }
    __syncthreads();
    for (int64_t r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
        const int p = pid[r];
        dst[r] = (int64_t)(base[p] + atomicAdd(&hist[p], 1u));
    }
}

template <typename T>
__global__ void __launch_bounds__(256) k_partition_scatter(int64_t rows, const int64_t* dst, const T* in, T* out) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) out[dst[r]] = in[r];
}

struct B24 {
    uint64_t a, b, c;
};
struct B32 {
    uint64_t a, b, c, d;
};

__global__ void __launch_bounds__(256) k_push_partial(PushArgs a) {
    char* slot = a.slot[blockIdx.x];
    unsigned long long n = *a.count;
    unsigned int err = *a.err;
    if (n > (unsigned long long)a.capRows) {  // more groups than a slot holds: not a "small result" (the caller repartitions instead)
        err |= 16u;                           // PGPU_ERR_OUTPUT_FULL
        n = 0;
    }
    for (int c = 0; c < a.ncols; c++) {
        const PushCol& pc = a.c[c];
        // arrays are 256-byte aligned at both ends and padded: whole 16-byte words
        const size_t vw = ((size_t)n * pc.stride + 15) / 16, nw = ((size_t)n + 15) / 16;
        const uint4* sv = (const uint4*)pc.values;
        uint4* dv = (uint4*)(slot + pc.dstV);
        for (size_t i = threadIdx.x; i < vw; i += blockDim.x) dv[i] = sv[i];
        const uint4* sn = (const uint4*)pc.nulls;
        uint4* dn = (uint4*)(slot + pc.dstN);
        for (size_t i = threadIdx.x; i < nw; i += blockDim.x) dn[i] = sn[i];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long* h = (unsigned long long*)slot;
        h[1] = n;
        h[2] = err;
        __threadfence_system();
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(h), "l"(a.seq) : "memory");
    }
}

int gridFor(int numSMs, int64_t rows) {
    int64_t g = (rows + 255) / 256;
    if (g < 1) g = 1;
    if (g > (int64_t)numSMs * 8) g = (int64_t)numSMs * 8;
    return (int)g;
}

}  // namespace

cudaError_t partitionCount(cudaStream_t stream, int numSMs, const PartKeys& keys, int64_t rows, uint8_t* pid, unsigned long long* counts) {
    k_partition_count<<<gridFor(numSMs, rows), 256, 0, stream>>>(keys, rows, pid, counts);
    return cudaGetLastError();
}

cudaError_t partitionAssign(cudaStream_t stream, int numSMs, int nparts, int64_t rows, const uint8_t* pid, unsigned long long* cursors, int64_t* dst) {
    k_partition_assign<<<gridFor(numSMs, rows), 256, 0, stream>>>(nparts, rows, pid, cursors, dst);
    return cudaGetLastError();
}

cudaError_t partitionScatter(cudaStream_t stream, int numSMs, int64_t rows, const int64_t* dst, const void* in, void* out, int width) {
    const int g = gridFor(numSMs, rows);
    switch (width) {
        case 1: k_partition_scatter<uint8_t><<<g, 256, 0, stream>>>(rows, dst, (const uint8_t*)in, (uint8_t*)out); break;
        case 4: k_partition_scatter<uint32_t><<<g, 256, 0, stream>>>(rows, dst, (const uint32_t*)in, (uint32_t*)out); break;
        case 8: k_partition_scatter<uint64_t><<<g, 256, 0, stream>>>(rows, dst, (const uint64_t*)in, (uint64_t*)out); break;
        case 16: k_partition_scatter<uint4><<<g, 256, 0, stream>>>(rows, dst, (const uint4*)in, (uint4*)out); break;
        case 24: k_partition_scatter<B24><<<g, 256, 0, stream>>>(rows, dst, (const B24*)in, (B24*)out); break;
        case 32: k_partition_scatter<B32><<<g, 256, 0, stream>>>(rows, dst, (const B32*)in, (B32*)out); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t pushPartial(cudaStream_t stream, const PushArgs& args) {
    k_push_partial<<<args.nranks, 256, 0, stream>>>(args);
    return cudaGetLastError();
}

int32_t partitionOfHost(const int64_t* keyBits, const uint8_t* isNull, const int32_t* kinds, int nkeys, int nparts) {
    int64_t h = 0;
    for (int k = 0; k < nkeys; k++) {
        int64_t hc = 0;
        if (!isNull[k]) hc = kinds[k] == PART_KEY_F64 ? hashDoubleBits(keyBits[k]) : kinds[k] == PART_KEY_I32 ? hashLong((int64_t)(int32_t)keyBits[k]) : hashLong(keyBits[k]);
        h = (int64_t)(31ull * (uint64_t)h + (uint64_t)hc);
    }
    return partitionOfHash(h, nparts);
}

}  // namespace pgpu
