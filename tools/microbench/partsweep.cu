// partsweep.cu — does ordering the rows by table region first pay for GROUP BY with 150 M groups?
// The table of randtable.cu (32-byte slots, 8.6 GB) takes 10-14 G random updates/s from HBM but 53 G/s when the slots touched
// stay L2 resident. Here: (A) count rows per partition (partition = top bits of the slot index), (B) scatter (key, value) rows
// into partition order — block-local stable counting sort staged in shared memory so that global writes are contiguous runs —
// (C) the same load / CAS / RED update loop over the reordered rows: every moment all blocks work inside one or two table
// slices of (table bytes / K).
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o partsweep partsweep.cu
// Run:   ./partsweep [rows=600000000] [clustered=0] [slots_log2=28]
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;
__device__ __forceinline__ u64 mix64(u64 x) { x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 27; x *= 0x94d049bb133111ebull; x ^= x >> 31; return x; }

__global__ void k_gen(u64* key, double* val, u64 n, u64 groups, int clustered) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        key[i] = 1 + (clustered ? i / 4 : mix64(i) % groups);
        val[i] = 1.0;
    }
}

// ---- (A) count
template <int KLOG>
__global__ void __launch_bounds__(256) k_count(const u64* __restrict__ key, u64 n, int slotsLog, u32* __restrict__ hist) {
    constexpr int K = 1 << KLOG;
    __shared__ u32 h[K];
    for (int j = threadIdx.x; j < K; j += 256) h[j] = 0;
    __syncthreads();
    for (u64 i = (u64)blockIdx.x * 256 + threadIdx.x; i < n; i += (u64)gridDim.x * 256) atomicAdd(&h[mix64(key[i]) >> (64 - KLOG)], 1u);
    __syncthreads();
    for (int j = threadIdx.x; j < K; j += 256)
        if (h[j]) atomicAdd(&hist[j], h[j]);
}
__global__ void k_scan(const u32* hist, u64* cursor, u64* start, int K) {
    u64 s = 0;
    for (int j = 0; j < K; j++) { cursor[j] = s; start[j] = s; s += hist[j]; }
    start[K] = s;
}

// ---- (B) scatter: tile of T rows per block; warp w owns rows [w*T/8, (w+1)*T/8) of the tile and walks them 32 at a time,
// so ranks are stable (row order is kept inside a partition => runs of equal keys survive)
template <int KLOG, int T>
__global__ void __launch_bounds__(256) k_scatter(const u64* __restrict__ key, const double* __restrict__ val, u64 n, u64* __restrict__ cursor,
                                                 u64* __restrict__ okey, double* __restrict__ oval) {
    constexpr int K = 1 << KLOG, W = 8, PER_WARP = T / W;
    extern __shared__ unsigned char smem[];
    u64* skey = (u64*)smem;                       // T
    double* sval = (double*)(skey + T);           // T
    unsigned short* whist = (unsigned short*)(sval + T);  // W * K: per-warp counts, then per-warp bases
    u32* lbase = (u32*)(whist + W * K);           // K + 1: block-local start of each partition
    u64* gbase = (u64*)(lbase + K + 2);           // K: global start of this tile's rows of each partition
    unsigned short* spid = (unsigned short*)(gbase + K);  // T: partition of the staged row
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const u64 ntiles = (n + T - 1) / T;
    for (u64 tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const u64 base = tile * T;
        for (int j = threadIdx.x; j < W * K; j += 256) whist[j] = 0;
        __syncthreads();
        // phase 1: per-warp counts
        for (int r = 0; r < PER_WARP; r += 32) {
            const u64 i = base + warp * PER_WARP + r + lane;
            const bool act = i < n;
            const u32 pid = act ? (u32)(mix64(key[i]) >> (64 - KLOG)) : 0xffffffffu;
            const u32 same = __match_any_sync(0xffffffffu, pid);
            if (act && lane == __ffs(same) - 1) whist[warp * K + pid] += (unsigned short)__popc(same);
        }
        __syncthreads();
        // block scan over partitions (thread j handles partitions j, j+256, ...): totals -> lbase, per-warp bases in place
        // first: total per partition
        for (int j = threadIdx.x; j < K; j += 256) {
            u32 t = 0;
            for (int w = 0; w < W; w++) t += whist[w * K + j];
            lbase[j] = t;
        }
        __syncthreads();
        if (warp == 0) {  // exclusive scan of K totals by one warp
            u32 carry = 0;
            for (int j0 = 0; j0 < K; j0 += 32) {
                u32 v = lbase[j0 + lane], x = v;
                for (int d = 1; d < 32; d <<= 1) { u32 y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
                lbase[j0 + lane] = carry + x - v;
                carry += __shfl_sync(0xffffffffu, x, 31);
            }
            if (lane == 0) lbase[K] = carry;
        }
        __syncthreads();
        for (int j = threadIdx.x; j < K; j += 256) {
            u32 b = lbase[j];
            const u32 tot = lbase[j + 1] - b;
            for (int w = 0; w < W; w++) { u32 c = whist[w * K + j]; whist[w * K + j] = (unsigned short)b; b += c; }
            gbase[j] = tot ? atomicAdd(&cursor[j], (u64)tot) : 0;
        }
        __syncthreads();
        // phase 2: stage rows in partition order
        for (int r = 0; r < PER_WARP; r += 32) {
            const u64 i = base + warp * PER_WARP + r + lane;
            const bool act = i < n;
            const u64 kx = act ? key[i] : 0;
            const u32 pid = act ? (u32)(mix64(kx) >> (64 - KLOG)) : 0xffffffffu;
            const u32 same = __match_any_sync(0xffffffffu, pid);
            if (act) {
                const u32 pos = whist[warp * K + pid] + __popc(same & ((1u << lane) - 1));
                skey[pos] = kx;
                sval[pos] = val[i];
                spid[pos] = (unsigned short)pid;
            }
            __syncwarp();
            if (act && lane == __ffs(same) - 1) whist[warp * K + pid] += (unsigned short)__popc(same);
            __syncwarp();
        }
        __syncthreads();
        const u32 cnt = lbase[K];
        for (u32 j = threadIdx.x; j < cnt; j += 256) {
            const u32 pid = spid[j];
            const u64 dst = gbase[pid] + (j - lbase[pid]);
            okey[dst] = skey[j];
            oval[dst] = sval[j];
        }
        __syncthreads();
    }
}

// ---- (C) update: thread takes R consecutive rows, merges neighbours with equal keys (a stand-in for run compaction), then per run:
// load the slot's key word, CAS when empty, linear probe on a different key, two REDs
template <int R>
__global__ void __launch_bounds__(256) k_update(const u64* __restrict__ key, const double* __restrict__ val, u64 n, u64* table, int slotsLog, u64* groups) {
    const u64 mask = (1ull << slotsLog) - 1;
    const u64 ntiles = (n + 256 * R - 1) / (256 * R);
    u64 fresh = 0;
    for (u64 tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const u64 i0 = tile * 256 * R + (u64)threadIdx.x * R;
        u64 k[R];
        double v[R];
#pragma unroll
        for (int r = 0; r < R; r++) {
            k[r] = i0 + r < n ? key[i0 + r] : 0;
            v[r] = i0 + r < n ? val[i0 + r] : 0.0;
        }
        u64 cnt[R];
#pragma unroll
        for (int r = 0; r < R; r++) cnt[r] = k[r] ? 1 : 0;
#pragma unroll
        for (int r = R - 1; r > 0; r--)
            if (k[r] && k[r] == k[r - 1]) { v[r - 1] += v[r]; cnt[r - 1] += cnt[r]; k[r] = 0; }
        u64 slot[R];
#pragma unroll
        for (int r = 0; r < R; r++) slot[r] = mix64(k[r]) >> (64 - slotsLog);
        u64 cur[R];
#pragma unroll
        for (int r = 0; r < R; r++)
            if (k[r]) cur[r] = *(volatile u64*)(table + slot[r] * 4);
#pragma unroll
        for (int r = 0; r < R; r++) {
            if (!k[r]) continue;
            u64 s = slot[r], c = cur[r];
            for (;;) {
                if (c == 0) {
                    c = atomicCAS(table + s * 4, 0ull, k[r]);
                    if (c == 0) { fresh++; c = k[r]; }
                }
                if (c == k[r]) break;
                s = (s + 1) & mask;
                c = *(volatile u64*)(table + s * 4);
            }
            atomicAdd((double*)(table + s * 4 + 2), v[r]);
            atomicAdd(table + s * 4 + 3, cnt[r]);
        }
    }
    if (fresh) atomicAdd(groups, fresh);
}

// ---- (C') the same update, and every tile also pulls the table lines that the rows `la` rows further on will touch into L2 with
// full-line sequential reads: without it every first touch of a slot is a 32-byte random DRAM read of its own
template <int R, int KLOG>
__global__ void __launch_bounds__(256) k_update_pf(const u64* __restrict__ key, const double* __restrict__ val, u64 n, u64* table, int slotsLog, u64* groups,
                                                   const u64* __restrict__ start, u64 la, u64* ticket) {
    constexpr int K = 1 << KLOG;
    const u64 mask = (1ull << slotsLog) - 1;
    const u64 ntiles = (n + 256 * R - 1) / (256 * R);
    const u64 sliceBytes = (32ull << slotsLog) >> KLOG;
    u64 fresh = 0;
    __shared__ u64 pf0, pf1, stile;
    for (;;) {
        if (threadIdx.x == 0) stile = ticket ? atomicAdd(ticket, 1ull) : 0;
        __syncthreads();
        const u64 tile = stile;
        if (tile >= ntiles) break;
        if (threadIdx.x == 0) {
            // table bytes that rows [x, x + 256 R) map to, x = first row of the tile + la
            u64 x[2] = {tile * 256 * R + la, tile * 256 * R + la + 256 * R};
            u64 b[2];
            for (int e = 0; e < 2; e++) {
                if (x[e] >= n) { b[e] = 32ull << slotsLog; continue; }
                int lo = 0, hi = K;  // last p with start[p] <= x
                while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (start[mid] <= x[e]) lo = mid; else hi = mid; }
                const u64 s0 = start[lo], s1 = start[lo + 1];
                b[e] = (u64)lo * sliceBytes + (u64)((double)(x[e] - s0) / (double)(s1 - s0 ? s1 - s0 : 1) * (double)sliceBytes);
            }
            pf0 = b[0] & ~127ull;
            pf1 = (b[1] + 127) & ~127ull;
        }
        __syncthreads();
        for (u64 off = pf0 + (u64)threadIdx.x * 128; off < pf1; off += 256 * 128)
            asm volatile("prefetch.global.L2 [%0];" ::"l"((const char*)table + off));
        const u64 i0 = tile * 256 * R + (u64)threadIdx.x * R;
        u64 k[R];
        double v[R];
#pragma unroll
        for (int r = 0; r < R; r++) {
            k[r] = i0 + r < n ? __ldcs(key + i0 + r) : 0;
            v[r] = i0 + r < n ? __ldcs(val + i0 + r) : 0.0;
        }
        u64 cnt[R];
#pragma unroll
        for (int r = 0; r < R; r++) cnt[r] = k[r] ? 1 : 0;
#pragma unroll
        for (int r = R - 1; r > 0; r--)
            if (k[r] && k[r] == k[r - 1]) { v[r - 1] += v[r]; cnt[r - 1] += cnt[r]; k[r] = 0; }
        u64 slot[R], cur[R];
#pragma unroll
        for (int r = 0; r < R; r++) slot[r] = mix64(k[r]) >> (64 - slotsLog);
#pragma unroll
        for (int r = 0; r < R; r++)
            if (k[r]) cur[r] = *(volatile u64*)(table + slot[r] * 4);
#pragma unroll
        for (int r = 0; r < R; r++)
            if (k[r] && cur[r] == 0) {
                cur[r] = atomicCAS(table + slot[r] * 4, 0ull, k[r]);
                if (cur[r] == 0) { fresh++; cur[r] = k[r]; }
            }
#pragma unroll
        for (int r = 0; r < R; r++) {
            if (!k[r]) continue;
            u64 s = slot[r], c = cur[r];
            while (c != k[r]) {
                s = (s + 1) & mask;
                c = *(volatile u64*)(table + s * 4);
                if (c == 0) {
                    c = atomicCAS(table + s * 4, 0ull, k[r]);
                    if (c == 0) { fresh++; c = k[r]; }
                }
            }
            atomicAdd((double*)(table + s * 4 + 2), v[r]);
            atomicAdd(table + s * 4 + 3, cnt[r]);
        }
        __syncthreads();
    }
    if (fresh) atomicAdd(groups, fresh);
}

// ---- (B') scatter, ranks from shared-memory atomics (row order inside a partition is not kept), only row indices staged
template <int KLOG, int T, int NT>
__global__ void __launch_bounds__(NT) k_scatter_u(const u64* __restrict__ key, const double* __restrict__ val, u64 n, u64* __restrict__ cursor,
                                                  u64* __restrict__ okey, double* __restrict__ oval) {
    constexpr int K = 1 << KLOG;
    extern __shared__ unsigned char smem[];
    u32* cnt = (u32*)smem;                         // K + 1: counts, then block-local starts
    u32* fill = cnt + K + 1;                       // K: running fill
    u64* gbase = (u64*)(fill + K + (K & 1 ? 0 : 1));  // K
    unsigned short* sidx = (unsigned short*)(gbase + K);  // T: tile-local row at sorted position
    unsigned short* spid = sidx + T;               // T: partition of the row at sorted position
    const int lane = threadIdx.x & 31;
    const u64 ntiles = (n + T - 1) / T;
    for (u64 tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const u64 base = tile * T;
        for (int j = threadIdx.x; j <= K; j += NT) cnt[j] = 0;
        __syncthreads();
        u32 pid[T / NT];
#pragma unroll
        for (int r = 0; r < T / NT; r++) {
            const u64 i = base + r * NT + threadIdx.x;
            pid[r] = i < n ? (u32)(mix64(key[i]) >> (64 - KLOG)) : 0xffffffffu;
            if (i < n) atomicAdd(&cnt[pid[r]], 1u);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            u32 carry = 0;
            for (int j0 = 0; j0 < K; j0 += 32) {
                u32 v = cnt[j0 + lane], x = v;
                for (int d = 1; d < 32; d <<= 1) { u32 y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
                cnt[j0 + lane] = carry + x - v;
                fill[j0 + lane] = carry + x - v;
                carry += __shfl_sync(0xffffffffu, x, 31);
            }
            if (lane == 0) cnt[K] = carry;
        }
        __syncthreads();
        for (int j = threadIdx.x; j < K; j += NT) {
            const u32 tot = cnt[j + 1] - cnt[j];
            gbase[j] = tot ? atomicAdd(&cursor[j], (u64)tot) : 0;
        }
#pragma unroll
        for (int r = 0; r < T / NT; r++) {
            if (pid[r] != 0xffffffffu) {
                const u32 pos = atomicAdd(&fill[pid[r]], 1u);
                sidx[pos] = (unsigned short)(r * NT + threadIdx.x);
                spid[pos] = (unsigned short)pid[r];
            }
        }
        __syncthreads();
        const u32 total = cnt[K];
        for (u32 j = threadIdx.x; j < total; j += NT) {
            const u32 p = spid[j];
            const u64 src = base + sidx[j], dst = gbase[p] + (j - cnt[p]);
            okey[dst] = key[src];
            oval[dst] = val[src];
        }
        __syncthreads();
    }
}

static float elapsed(cudaEvent_t a, cudaEvent_t b) { float ms; cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b); return ms; }
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

template <int KLOG, int T>
void sweep(const u64* key, const double* val, u64* okey, double* oval, u64 n, u64* table, int slotsLog, u32* hist, u64* cursor, u64* start, u64* groups, int updBlocks) {
    constexpr int K = 1 << KLOG;
    const size_t smem = (size_t)T * 16 + (size_t)8 * K * 2 + (K + 2) * 4 + (size_t)K * 8 + (size_t)T * 2 + 16;
    CK(cudaFuncSetAttribute(k_scatter<KLOG, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e[5];
    for (auto& x : e) cudaEventCreate(&x);
    CK(cudaMemset(table, 0, (size_t)32 << slotsLog));
    CK(cudaMemset(hist, 0, K * 4));
    CK(cudaMemset(groups, 0, 8));
    cudaEventRecord(e[0]);
    k_count<KLOG><<<148 * 8, 256>>>(key, n, slotsLog, hist);
    k_scan<<<1, 1>>>(hist, cursor, start, K);
    cudaEventRecord(e[1]);
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_scatter<KLOG, T>, 256, smem));
    k_scatter<KLOG, T><<<148 * (occ > 0 ? occ : 1), 256, smem>>>(key, val, n, cursor, okey, oval);
    cudaEventRecord(e[2]);
    k_update<4><<<148 * updBlocks, 256>>>(okey, oval, n, table, slotsLog, groups);
    cudaEventRecord(e[3]);
    CK(cudaDeviceSynchronize());
    u64 g = 0;
    CK(cudaMemcpy(&g, groups, 8, cudaMemcpyDeviceToHost));
    const float a = elapsed(e[0], e[1]), b = elapsed(e[1], e[2]), c = elapsed(e[2], e[3]);
    printf("K=%4d T=%5d scatter blocks/SM=%d update blocks/SM=%d: count %.2f ms, scatter %.2f ms (%.0f GB/s), update %.2f ms (%.1f G rows/s), total %.2f ms; groups %llu\n",
           K, T, occ, updBlocks, a, b, n * 32.0 / b / 1e6, c, n / c / 1e6, a + b + c, g);
}

template <int KLOG, int T, int NT>
void sweep2(const u64* key, const double* val, u64* okey, double* oval, u64 n, u64* table, int slotsLog, u32* hist, u64* cursor, u64* start, u64* groups, u64 la) {
    constexpr int K = 1 << KLOG;
    const size_t smem = (size_t)(2 * K + 2) * 4 + (size_t)K * 8 + (size_t)T * 4 + 16;
    CK(cudaFuncSetAttribute(k_scatter_u<KLOG, T, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e[5];
    for (auto& x : e) cudaEventCreate(&x);
    CK(cudaMemset(table, 0, (size_t)32 << slotsLog));
    CK(cudaMemset(hist, 0, K * 4));
    CK(cudaMemset(groups, 0, 8));
    cudaEventRecord(e[0]);
    k_count<KLOG><<<148 * 8, 256>>>(key, n, slotsLog, hist);
    k_scan<<<1, 1>>>(hist, cursor, start, K);
    cudaEventRecord(e[1]);
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_scatter_u<KLOG, T, NT>, NT, smem));
    k_scatter_u<KLOG, T, NT><<<148 * (occ > 0 ? occ : 1), NT, smem>>>(key, val, n, cursor, okey, oval);
    cudaEventRecord(e[2]);
    CK(cudaMemset(groups + 1, 0, 8));
    if (la) k_update_pf<4, KLOG><<<148 * 8, 256>>>(okey, oval, n, table, slotsLog, groups, start, la, groups + 1);
    else k_update<4><<<148 * 8, 256>>>(okey, oval, n, table, slotsLog, groups);
    cudaEventRecord(e[3]);
    CK(cudaDeviceSynchronize());
    u64 g = 0;
    CK(cudaMemcpy(&g, groups, 8, cudaMemcpyDeviceToHost));
    const float a = elapsed(e[0], e[1]), b = elapsed(e[1], e[2]), c = elapsed(e[2], e[3]);
    printf("v2 K=%4d T=%5d NT=%d scatter blocks/SM=%d lookahead %llu rows: count %.2f ms, scatter %.2f ms (%.0f GB/s), update %.2f ms (%.1f G rows/s), total %.2f ms; groups %llu\n",
           K, T, NT, occ, la, a, b, n * 32.0 / b / 1e6, c, n / c / 1e6, a + b + c, g);
}

int main(int argc, char** argv) {
    const u64 n = argc > 1 ? strtoull(argv[1], 0, 10) : 600000000ull;
    const int clustered = argc > 2 ? atoi(argv[2]) : 0;
    const int slotsLog = argc > 3 ? atoi(argv[3]) : 28;
    u64 *key, *okey, *table, *cursor, *start, *groups;
    double *val, *oval;
    u32* hist;
    CK(cudaMalloc(&key, n * 8));
    CK(cudaMalloc(&val, n * 8));
    CK(cudaMalloc(&okey, n * 8));
    CK(cudaMalloc(&oval, n * 8));
    CK(cudaMalloc(&table, (size_t)32 << slotsLog));
    CK(cudaMalloc(&hist, 4096 * 4));
    CK(cudaMalloc(&cursor, 4097 * 8));
    CK(cudaMalloc(&start, 4097 * 8));
    CK(cudaMalloc(&groups, 16));
    k_gen<<<148 * 8, 256>>>(key, val, n, n / 4, clustered);
    CK(cudaDeviceSynchronize());
    printf("%llu rows, %s keys, table 2^%d slots x 32 B = %.1f GB\n", n, clustered ? "clustered (runs of 4)" : "shuffled", slotsLog, (32ull << slotsLog) / 1e9);
    // baseline: the update loop over the rows as they come
    for (int rep = 0; rep < 2; rep++) {
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        CK(cudaMemset(table, 0, (size_t)32 << slotsLog));
        CK(cudaMemset(groups, 0, 8));
        cudaEventRecord(a);
        k_update<4><<<148 * 8, 256>>>(key, val, n, table, slotsLog, groups);
        cudaEventRecord(b);
        const float ms = elapsed(a, b);
        u64 g = 0;
        CK(cudaMemcpy(&g, groups, 8, cudaMemcpyDeviceToHost));
        printf("unpartitioned update: %.2f ms (%.1f G rows/s); groups %llu\n", ms, n / ms / 1e6, g);
    }
    if (getenv("ONLY")) {
        sweep2<10, 8192, 1024>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, getenv("LA") ? atoll(getenv("LA")) : 3000000);
        return 0;
    }
    sweep<9, 4096>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 8);
    sweep<10, 4096>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 8);
    for (u64 la : {300000ull, 600000ull, 1000000ull, 1500000ull, 2500000ull}) sweep2<10, 8192, 1024>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, la);
    for (u64 la : {300000ull, 600000ull, 1000000ull, 1500000ull, 2500000ull}) sweep2<9, 8192, 1024>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, la);
    for (u64 la : {600000ull, 1500000ull}) sweep2<11, 8192, 1024>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, la);
    for (u64 la : {600000ull, 1500000ull, 4000000ull}) sweep2<8, 8192, 1024>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, la);
    return 0;
    sweep2<9, 4096, 256>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 0);
    sweep2<9, 4096, 256>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 2000000);
    sweep2<9, 4096, 256>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 4000000);
    sweep2<10, 4096, 256>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 1500000);
    sweep2<10, 4096, 256>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 3000000);
    sweep2<10, 4096, 256>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 6000000);
    sweep2<10, 8192, 512>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 3000000);
    sweep2<10, 8192, 1024>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 3000000);
    sweep2<11, 8192, 512>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 3000000);
    sweep2<11, 4096, 256>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 3000000);
    sweep2<8, 4096, 256>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 4000000);
    sweep2<7, 4096, 256>(key, val, okey, oval, n, table, slotsLog, hist, cursor, start, groups, 4000000);
    return 0;
}
