// randtable.cu — what HBM3e gives a hash table whose every update is a random 32-byte slot: the floor under GROUP BY with
// shuffled keys (BASELINE configs[2], shuffled variant). 600 M updates into an 8 GB table (far larger than the 126 MB L2):
//   mode 0: one RED.ADD.F64 per update                         (fire-and-forget, one sector read-modify-write)
//   mode 1: 16-byte load of the slot's key words, then two REDs (what a table hit costs)
//   mode 2: load, compare-and-swap of the key words, two REDs   (what a new group costs)
//   mode 3: mode 1 with a prefetch (L2 only) issued one tile ahead
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o randtable randtable.cu ; run: ./randtable [slots_log2=28] [updates=600000000]
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 mix64(u64 x) { x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 27; x *= 0x94d049bb133111ebull; x ^= x >> 31; return x; }
template <int MODE, int R>
__global__ void __launch_bounds__(256) k(u64* table, u64 mask, u64 n) {
    const u64 stride = (u64)gridDim.x * blockDim.x * R;
    for (u64 i = ((u64)blockIdx.x * blockDim.x + threadIdx.x) * R; i < n; i += stride) {
        u64* s[R];
#pragma unroll
        for (int r = 0; r < R; r++) s[r] = table + (mix64(i + r) & mask) * 4;
        if (MODE == 3) {
#pragma unroll
            for (int r = 0; r < R; r++) {
                const u64* nx = table + (mix64(i + stride + r) & mask) * 4;
                asm volatile("prefetch.global.L2 [%0];" ::"l"(nx));
            }
        }
        u64 c0[R], c1[R];
        if (MODE >= 1) {
#pragma unroll
            for (int r = 0; r < R; r++)
                asm volatile("{\n .reg .b128 v;\n ld.relaxed.gpu.global.b128 v, [%2];\n mov.b128 {%0, %1}, v;\n}" : "=l"(c0[r]), "=l"(c1[r]) : "l"(s[r]) : "memory");
        }
        if (MODE == 2) {
#pragma unroll
            for (int r = 0; r < R; r++) {
                if (c1[r] == 0) {
                    u64 o0, o1;
                    asm volatile("{\n .reg .b128 c, n, o;\n mov.b128 c, {%2, %3};\n mov.b128 n, {%4, %5};\n atom.relaxed.gpu.global.cas.b128 o, [%6], c, n;\n mov.b128 {%0, %1}, o;\n}"
                                 : "=l"(o0), "=l"(o1) : "l"(0ull), "l"(0ull), "l"(i + r), "l"(0x8000000000000000ull), "l"(s[r]) : "memory");
                    c1[r] = o1;
                }
            }
        }
#pragma unroll
        for (int r = 0; r < R; r++) {
            if (MODE == 0 || c1[r] != 1) {
                atomicAdd((double*)(s[r] + 2), 1.0);
                if (MODE >= 1) atomicAdd(s[r] + 3, 1ull);
            }
        }
    }
}
template <int MODE, int R>
void run(const char* name, u64* table, u64 slots, u64 n, int blocksPerSM) {
    cudaMemset(table, 0, slots * 32);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a);
    k<MODE, R><<<148 * blocksPerSM, 256>>>(table, slots - 1, n);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    printf("%-40s R=%d blocks/SM=%d: %.2f ms, %.2f G updates/s  (%s)\n", name, R, blocksPerSM, ms, n / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}
int main(int argc, char** argv) {
    const int lg = argc > 1 ? atoi(argv[1]) : 28;
    const u64 slots = 1ull << lg, n = argc > 2 ? strtoull(argv[2], 0, 10) : 600000000ull;
    u64* table;
    if (cudaMalloc(&table, slots * 32) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    printf("table: %llu slots x 32 B = %.1f GB, %llu updates\n", slots, slots * 32 / 1e9, n);
    run<0, 4>("RED only", table, slots, n, 8);
    run<0, 8>("RED only", table, slots, n, 8);
    run<1, 4>("load + 2 RED", table, slots, n, 8);
    run<1, 8>("load + 2 RED", table, slots, n, 8);
    run<1, 8>("load + 2 RED", table, slots, n, 4);
    run<1, 16>("load + 2 RED", table, slots, n, 4);
    run<2, 4>("load + CAS + 2 RED", table, slots, n, 8);
    run<2, 8>("load + CAS + 2 RED", table, slots, n, 8);
    run<3, 4>("prefetch.L2 ahead + load + 2 RED", table, slots, n, 8);
    run<3, 8>("prefetch.L2 ahead + load + 2 RED", table, slots, n, 8);
    // a table that fits L2 (64 MB): the partitioned alternative's inner loop
    run<1, 8>("load + 2 RED, 64 MB table (L2 resident)", table, 1ull << 21, n, 8);
    run<2, 8>("load + CAS + 2 RED, 64 MB table", table, 1ull << 21, n, 8);
    return 0;
}
