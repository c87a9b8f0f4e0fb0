// pgpu_device.cuh — hand-written sm_100a device library of the fused scan -> filter -> project -> aggregate
// pipeline. The plan compiler (codegen.cpp) emits a small traits struct `P` (column loads, the RowExpression
// subset as straight-line CUDA, key packing, accumulator contributions, output finalisation) and instantiates
// the kernel templates below with it; NVRTC compiles the result for sm_100a (jit.cpp).
//
// What the kernels replace in the reference (paths under core/trino-main/src/main/java/io/trino/):
//   k_agg_global   PageProcessor (operator/project/PageProcessor.java:111-137,307-347) + AggregationOperator.addInput
//                  (operator/AggregationOperator.java:145-161) fused: one pass, no SelectedPositions materialised.
//   k_agg_hash     PageProcessor + InMemoryHashAggregationBuilder.processPage (operator/aggregation/builder/
//                  InMemoryHashAggregationBuilder.java:139-155): GroupByHash.getGroupIds (operator/BigintGroupByHash.java:
//                  213-244, MultiChannelGroupByHash.java:281-342) and every GroupedAccumulator.addInput
//                  (operator/aggregation/AccumulatorCompiler.java:490-569) in one pass; no GroupByIdBlock.
//   k_compact      PageFilter.positionsArrayToSelectedPositions (operator/project/PageFilter.java:27-50) +
//                  PageProjection.project at the selected positions (sql/gen/PageFunctionCompiler.java:283-319).
//   k_materialize  InMemoryHashAggregationBuilder.buildResult (…/InMemoryHashAggregationBuilder.java:272-298).
//   k_rehash       BigintGroupByHash.tryRehash (operator/BigintGroupByHash.java:262-313).
//
// Memory-system design (HBM-bound integer/fp64 work, no tensor cores): every thread owns R consecutive rows so
// each fixed-width column is read with 128-bit, fully coalesced, L1-bypassing loads (ld.global.nc.L1::no_allocate);
// grids are persistent (a multiple of the SM count, host side) and stride over tiles; accumulators live in
// registers (global aggregation), in warp-private shared-memory tables keyed through a block-shared key directory
// (low cardinality), or in an open-addressing table in HBM updated with fire-and-forget RED atomics (high cardinality).
#pragma once

typedef long long i64;
typedef unsigned long long u64;
typedef int i32;
typedef unsigned int u32;
typedef unsigned char u8;
typedef signed char i8;
typedef __int128 i128;
typedef unsigned __int128 u128;

#define PGPU_DEV __device__ __forceinline__
#define FULLMASK 0xffffffffu

// sticky error bits (host maps them to pgpu_status / TrinoException codes)
#define PGPU_ERR_OVERFLOW 1u     // NUMERIC_VALUE_OUT_OF_RANGE
#define PGPU_ERR_DIVZERO 2u      // DIVISION_BY_ZERO
#define PGPU_ERR_TABLE_FULL 4u   // internal: probe chain exhausted (host sizes tables so this cannot happen)
#define PGPU_ERR_VARCHAR_LEN 8u  // a varchar(n) key longer than n bytes
#define PGPU_ERR_OUTPUT_FULL 16u // internal: output buffer too small
#define PGPU_ERR_INVALID_CAST 32u// INVALID_CAST_ARGUMENT

// One input Block as the kernels see it (device pointers). Plain blocks use values/nulls/offsets;
// dictionary blocks use ids + d*; RLE blocks use d* with a single position.
struct ColArg {
    const void* values;
    const u8* nulls;
    const i32* offsets;
    const i32* ids;
    const void* dvalues;
    const u8* dnulls;
    const i32* doffsets;
    i32 dcount;  // entries of the dictionary
    i32 pad;
};

template <int NC>
struct PageArgs {
    ColArg c[NC];
    i64 rows;
};

// Open-addressing table in HBM. Slot = [tag][key words][cell words], SLOTW 64-bit words, 32-byte aligned.
struct TableArgs {
    u64* slots;
    u64 capacity;    // slots
    u64* counters;   // [0] = groups inserted, [1] = rows deferred because the table reached `limit`
    u64 limit;       // new groups are only claimed while counters[0] < limit (load-factor guard)
    u64* faillog;    // chunk-relative indices of deferred rows; replayed by k_replay after the host grew the table
    u64 faillog_cap;
    u64 row_bias;    // added to the kernel's row numbers when logging (a kernel launched on the tail of a chunk)
    u64 bucket_off;  // bucketed tables: added to the key before it is cut into (bucket, position), so that clusters that start
                     // at 8n + 1 (TPC-H order keys: 1..8, 33..40, ...) fill one bucket instead of straddling two
};

// Per-pipeline device scratch shared by all kernels.
struct Scratch {
    u32* err;        // sticky error bits
    u64* partials;   // k_agg_global: [grid][NW] block partials
    u32* ticket;     // k_agg_global / k_compact: block tickets (zeroed by the finishing block)
    u64* state;      // k_agg_global: the single accumulator slot [NW]
    u64* tile_status;// k_compact: decoupled look-back descriptors
    u64* out_count;  // k_compact / k_materialize: rows written
};

struct OutCol {
    void* values;    // fixed width: element array; var-width: 16-byte cells (len byte + 15 data bytes)
    u8* nulls;       // 1 byte / position
};

template <int NO>
struct OutArgs {
    OutCol c[NO];
    i64 capacity;    // rows the buffers can hold
};

// ------------------------------------------------------------------------------------------------ loads
#ifdef PGPU_HOST_EMULATION
// tests/emu/expr_emu.py compiles this header's scalar part and a plan's generated P::eval_project as plain C++ to run the
// expression code on the CPU against the known answers; the kernels below the scalar section are not part of that build.
PGPU_DEV int4 ld_stream16(const void* p) { int4 r; memcpy(&r, p, 16); return r; }
PGPU_DEV u32 ld_stream4(const void* p) { u32 r; memcpy(&r, p, 4); return r; }
PGPU_DEV u64 ld_stream8(const void* p) { u64 r; memcpy(&r, p, 8); return r; }
#else
PGPU_DEV int4 ld_stream16(const void* p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
PGPU_DEV u32 ld_stream4(const void* p) {
    u32 r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
PGPU_DEV u64 ld_stream8(const void* p) {
    u64 r;
    asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(r) : "l"(p));
    return r;
}
#endif  // PGPU_HOST_EMULATION

// Load the R (=4) consecutive elements [base, base+4) of a flat column. `full` = the whole tile is in range and the
// column pointer is 16-byte aligned (host guarantees alignment; base is a multiple of 4).
template <typename T>
PGPU_DEV void load_tile4(const T* __restrict__ p, i64 base, i64 nrows, bool full, T (&v)[4]) {
    if (full) {
        if (sizeof(T) == 8) {
            int4 a = ld_stream16(p + base), b = ld_stream16(p + base + 2);
            u64 w[4];
            w[0] = ((u64)(u32)a.y << 32) | (u32)a.x;
            w[1] = ((u64)(u32)a.w << 32) | (u32)a.z;
            w[2] = ((u64)(u32)b.y << 32) | (u32)b.x;
            w[3] = ((u64)(u32)b.w << 32) | (u32)b.z;
#pragma unroll
            for (int r = 0; r < 4; r++) v[r] = *reinterpret_cast<T*>(&w[r]);
        } else if (sizeof(T) == 4) {
            int4 a = ld_stream16(p + base);
            u32 w[4] = {(u32)a.x, (u32)a.y, (u32)a.z, (u32)a.w};
#pragma unroll
            for (int r = 0; r < 4; r++) v[r] = *reinterpret_cast<T*>(&w[r]);
        } else if (sizeof(T) == 1) {
            u32 a = ld_stream4(p + base);
#pragma unroll
            for (int r = 0; r < 4; r++) {
                u8 b = (u8)(a >> (8 * r));
                v[r] = *reinterpret_cast<T*>(&b);
            }
        } else {  // 16-byte elements (Int128ArrayBlock)
#pragma unroll
            for (int r = 0; r < 4; r++) {
                int4 a = ld_stream16(p + base + r);
                v[r] = *reinterpret_cast<T*>(&a);
            }
        }
    } else {
#pragma unroll
        for (int r = 0; r < 4; r++) {
            if (base + r < nrows) v[r] = p[base + r];
            else v[r] = T();
        }
    }
}

struct raw128 {
    u64 lo, hi;
};

// Int128ArrayBlock element (sign-magnitude, low word first, sign = top bit of word 1;
// spi/type/UnscaledDecimal128Arithmetic.java:30-56) <-> two's complement.
PGPU_DEV i128 dec_from_raw(raw128 r) {
    u128 mag = ((u128)(r.hi & 0x7fffffffffffffffull) << 64) | r.lo;
    return (r.hi >> 63) ? -(i128)mag : (i128)mag;
}
PGPU_DEV raw128 dec_to_raw(i128 v) {
    bool neg = v < 0;
    u128 mag = neg ? (u128)0 - (u128)v : (u128)v;
    raw128 r;
    r.lo = (u64)mag;
    r.hi = (u64)(mag >> 64) | (neg ? 0x8000000000000000ull : 0ull);
    return r;
}
// |v| > 10^38 - 1  (Decimals.MAX_UNSCALED_DECIMAL; UnscaledDecimal128Arithmetic.throwIfOverflows)
PGPU_DEV bool dec_overflows(i128 v) {
    const u128 lim = ((u128)0x4b3b4ca85a86c47aull << 64) | 0x098a224000000000ull;  // 10^38
    u128 mag = v < 0 ? (u128)0 - (u128)v : (u128)v;
    return mag >= lim;
}
// exact 127-bit-magnitude product (UnscaledDecimal128Arithmetic.multiply :452-650: throws when the magnitude
// needs more than 127 bits); `check38` adds throwIfOverflows (|r| >= 10^38) as DecimalOperators.multiply* do.
PGPU_DEV i128 dec_mul_impl(i128 a, i128 b, bool check38, u32& err) {
    bool neg = (a < 0) != (b < 0);
    u128 ma = a < 0 ? (u128)0 - (u128)a : (u128)a, mb = b < 0 ? (u128)0 - (u128)b : (u128)b;
    u64 a0 = (u64)ma, a1 = (u64)(ma >> 64), b0 = (u64)mb, b1 = (u64)(mb >> 64);
    bool ovf = (a1 != 0 && b1 != 0);
    u128 lo = (u128)a0 * b0;
    u128 c1 = (u128)a0 * b1, c2 = (u128)a1 * b0;
    ovf |= (c1 >> 64) != 0 || (c2 >> 64) != 0;
    u128 mid = (u128)(u64)c1 + (u128)(u64)c2 + (lo >> 64);
    ovf |= (mid >> 63) != 0;
    u128 m = ((u128)(u64)mid << 64) | (u64)lo;
    i128 r = neg ? -(i128)m : (i128)m;
    if (ovf || (check38 && dec_overflows(r))) err |= PGPU_ERR_OVERFLOW;
    return r;
}
PGPU_DEV i128 dec_mul(i128 a, i128 b, u32& err) { return dec_mul_impl(a, b, true, err); }
// a * 10^k (UnscaledDecimal128Arithmetic.rescale :218-232, upscale direction: overflow only past 127 bits); p10 = 10^k
PGPU_DEV i128 dec_upscale(i128 a, i128 p10, u32& err) { return dec_mul_impl(a, p10, false, err); }
// HALF_UP division by 10^k (rescale down / avg): round half away from zero
PGPU_DEV i128 dec_div_round(i128 a, i128 d) {
    bool neg = (a < 0) != (d < 0);
    u128 ma = a < 0 ? (u128)0 - (u128)a : (u128)a, md = d < 0 ? (u128)0 - (u128)d : (u128)d;
    u128 q = ma / md, rem = ma % md;
    if (rem >= md - rem) q += 1;  // 2*rem >= md without overflow
    return neg ? -(i128)q : (i128)q;
}

// ------------------------------------------------------------------------------------------------ scalars
PGPU_DEV u64 d2u(double d) { return (u64)__double_as_longlong(d); }
PGPU_DEV double u2d(u64 u) { return __longlong_as_double((i64)u); }
PGPU_DEV i64 add_i64_checked(i64 a, i64 b, u32& err) {  // Math.addExact (main/type/BigintOperators.java:47-56)
    i64 r = (i64)((u64)a + (u64)b);
    if (((a ^ r) & (b ^ r)) < 0) err |= PGPU_ERR_OVERFLOW;
    return r;
}
PGPU_DEV i64 sub_i64_checked(i64 a, i64 b, u32& err) {
    i64 r = (i64)((u64)a - (u64)b);
    if (((a ^ b) & (a ^ r)) < 0) err |= PGPU_ERR_OVERFLOW;
    return r;
}
PGPU_DEV i64 mul_i64_checked(i64 a, i64 b, u32& err) {
    i64 lo = (i64)((u64)a * (u64)b);
    i64 hi = __mul64hi(a, b);
    if (hi != (lo >> 63)) err |= PGPU_ERR_OVERFLOW;
    return lo;
}
PGPU_DEV i64 div_i64_checked(i64 a, i64 b, u32& err) {  // BigintOperators.divide :84-96
    if (b == 0) { err |= PGPU_ERR_DIVZERO; return 0; }
    if (a == (i64)0x8000000000000000ull && b == -1) { err |= PGPU_ERR_OVERFLOW; return 0; }
    return a / b;
}
PGPU_DEV i64 mod_i64_checked(i64 a, i64 b, u32& err) {  // BigintOperators.modulus :98-108 (Java: MIN % -1 == 0)
    if (b == 0) { err |= PGPU_ERR_DIVZERO; return 0; }
    if (b == -1) return 0;
    return a % b;
}
PGPU_DEV i64 neg_i64_checked(i64 a, u32& err) {
    if (a == (i64)0x8000000000000000ull) err |= PGPU_ERR_OVERFLOW;
    return (i64)(0 - (u64)a);
}
PGPU_DEV i32 narrow_i32_checked(i64 v, u32& err) {  // Math.toIntExact / IntegerOperators addExact family
    if (v != (i64)(i32)v) err |= PGPU_ERR_OVERFLOW;
    return (i32)v;
}
// IntegerOperators.divide (main/type/IntegerOperators.java:79-89) divides Java longs: MIN / -1 is 2^31, which leaves the INTEGER
// domain and fails when it is written to a block (AbstractIntType.writeLong: "Value 2147483648 exceeds MAX_INT"). Here it fails
// at the division, as an out-of-range value; it never wraps silently.
PGPU_DEV i32 div_i32_checked(i32 a, i32 b, u32& err) {
    if (b == 0) { err |= PGPU_ERR_DIVZERO; return 0; }
    if (b == -1) {
        if (a == (i32)0x80000000u) err |= PGPU_ERR_OVERFLOW;
        return (i32)(0u - (u32)a);
    }
    return a / b;
}
PGPU_DEV i32 mod_i32_checked(i32 a, i32 b, u32& err) {
    if (b == 0) { err |= PGPU_ERR_DIVZERO; return 0; }
    if (b == -1) return 0;
    return a % b;
}
// DoubleOperators.castToLong (main/type/DoubleOperators.java:155-165): Guava DoubleMath.roundToLong(value, HALF_UP)
// = round half away from zero; NaN, infinities and values outside [-2^63, 2^63) are INVALID_CAST_ARGUMENT.
PGPU_DEV i64 cast_f64_i64(double v, u32& err) {
    if (!(v == v) || fabs(v) == __longlong_as_double(0x7ff0000000000000ll)) { err |= PGPU_ERR_INVALID_CAST; return 0; }
    double r = trunc(v);
    if (fabs(v - r) >= 0.5) r += (v < 0.0 ? -1.0 : 1.0);
    if (!(r >= -9223372036854775808.0 && r < 9223372036854775808.0)) { err |= PGPU_ERR_INVALID_CAST; return 0; }
    return (i64)r;
}
// Double.compare total order as a signed integer (NaNs canonicalised, -0.0 < +0.0)
PGPU_DEV i64 f64_order_key(double v) {
    if (v != v) return 0x7ff8000000000000ll;
    i64 b = __double_as_longlong(v);
    return b ^ ((b >> 63) & 0x7fffffffffffffffll);
}
PGPU_DEV double f64_from_order_key(i64 k) { return __longlong_as_double(k ^ ((k >> 63) & 0x7fffffffffffffffll)); }
// grouping canonical form of a DOUBLE key (IS DISTINCT FROM semantics, spi/type/DoubleType.java:163-195)
PGPU_DEV u64 f64_group_bits(double v) {
    if (v != v) return 0x7ff8000000000000ull;
    if (v == 0.0) return 0ull;
    return d2u(v);
}

// Pack a var-width value of at most N (<=15) bytes into 16 bytes: byte 0 = length, bytes 1.. = data.
template <int N>
PGPU_DEV void pack_varchar(const u8* __restrict__ bytes, i32 off, i32 len, u64& lo, u64& hi, u32& err) {
    lo = 0; hi = 0;
    if (len > N) { err |= PGPU_ERR_VARCHAR_LEN; len = N; }
    if (len < 0) len = 0;
    lo = (u64)(u32)len;
#pragma unroll
    for (int k = 0; k < N; k++) {
        if (k < len) {
            u64 b = bytes[off + k];
            if (k < 7) lo |= b << (8 * (k + 1));
            else hi |= b << (8 * (k - 7));
        }
    }
}

// The <= MAXB (<= 16) bytes [start, start + nbytes) of a byte slice as 128 bits (little endian), fetched with aligned
// 32-bit loads: the var-width payload of a thread's 4 consecutive short rows is one contiguous window of the slice.
// MAXB = 4 * n for varchar(n): short windows need fewer words (varchar(1): two loads and one funnel shift).
template <int MAXB>
PGPU_DEV void load_byte_window(const u8* __restrict__ bytes, i32 start, i32 nbytes, u64& lo, u64& hi, u32& err) {
    lo = 0; hi = 0;
    if (nbytes <= 0) return;
    if (nbytes > MAXB) { err |= PGPU_ERR_VARCHAR_LEN; nbytes = MAXB; }
    const u32* wp = reinterpret_cast<const u32*>(bytes + (start & ~3));
    const u32 sh = (u32)(start & 3) * 8u;
    const int nwords = ((start & 3) + nbytes + 3) >> 2;  // 1 .. (MAXB + 6) / 4
    constexpr int MAXW = (MAXB + 6) / 4;
    u32 w[5] = {0u, 0u, 0u, 0u, 0u};
#pragma unroll
    for (int k = 0; k < MAXW; k++) w[k] = k < nwords ? __ldg(wp + k) : 0u;
    u32 o0 = __funnelshift_r(w[0], w[1], sh);
    u32 o1 = MAXB > 4 ? __funnelshift_r(w[1], w[2], sh) : 0u;
    lo = ((u64)o1 << 32) | o0;
    if (MAXB > 8) {
        u32 o2 = __funnelshift_r(w[2], w[3], sh), o3 = __funnelshift_r(w[3], w[4], sh);
        hi = ((u64)o3 << 32) | o2;
    }
}
// Row value out of a window of at most WB bytes: bytes [rel, rel + len) -> packed cell (byte 0 = length, bytes 1.. =
// data), N <= 4
template <int N, int WB>
PGPU_DEV void window_varchar(u64 wlo, u64 whi, i32 rel, i32 len, u64& x, u32& err) {
    if (len > N) { err |= PGPU_ERR_VARCHAR_LEN; len = N; }
    if (len < 0) len = 0;
    u64 d;
    if (WB <= 8) {
        d = wlo >> (8 * (rel & 7));  // the whole window lives in the low word
    } else {
        if (rel < 8) d = (wlo >> (8 * rel)) | (rel ? (whi << (64 - 8 * rel)) : 0ull);
        else d = whi >> (8 * (rel - 8));
    }
    d &= (1ull << (8 * len)) - 1ull;  // len <= N <= 4
    x = (u64)(u32)len | (d << 8);
}

#ifndef PGPU_HOST_EMULATION  // everything below needs a GPU: warp operations, atomics, shared memory, the kernels
// ------------------------------------------------------------------------------------------------ hashing / warp ops
PGPU_DEV u64 mix64(u64 x) {  // splitmix64 finaliser; table placement only, never observable
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}
// Table placement. PGPU_LOCB > 0 (set by the plan compiler for high-cardinality plans) keeps the 2^LOCB keys that differ only
// in the low bits of the first key field (bit PGPU_LOCSH of key word 0 upwards) in 2^LOCB CONSECUTIVE home slots: the
// neighbourhood is hashed, the low bits index inside it. Clustered keys (l_orderkey runs of consecutive rows) then share
// DRAM sectors and pages instead of taking one random sector each; uniformly random keys are placed as before.
#ifndef PGPU_LOCB
#define PGPU_LOCB 0
#endif
#ifndef PGPU_LOCSH
#define PGPU_LOCSH 0
#endif
template <int KW>
PGPU_DEV u64 hash_key(const u64 (&k)[KW > 0 ? KW : 1]) {
    u64 h = 0x9e3779b97f4a7c15ull;
    constexpr u64 LOCM = (1ull << PGPU_LOCB) - 1ull;
#pragma unroll
    for (int i = 0; i < KW; i++) h = mix64(h ^ (i == 0 ? k[i] & ~(LOCM << PGPU_LOCSH) : k[i]));
    if (PGPU_LOCB > 0) h = (h & ~LOCM) | (((k[0] >> PGPU_LOCSH) + (h >> 32)) & LOCM);  // rotated inside the neighbourhood block
    return h;
}
// home slot of a hash value
PGPU_DEV u64 home_pos(u64 h, u64 capacity) {
    if (PGPU_LOCB == 0) return __umul64hi(h, capacity);
    return (__umul64hi(h, capacity >> PGPU_LOCB) << PGPU_LOCB) | (h & ((1ull << PGPU_LOCB) - 1ull));
}

// cell-word kinds
#define CK_ADD_U64 0
#define CK_ADD_F64 1
#define CK_MAX_U64 2
#define CK_CNT 3  // ADD_U64 whose per-lane contribution is a small row count (< 2^26): warp-reduced with one REDUX

template <int KIND>
PGPU_DEV u64 cell_combine(u64 a, u64 b) {
    if (KIND == CK_ADD_U64 || KIND == CK_CNT) return a + b;
    if (KIND == CK_ADD_F64) return d2u(u2d(a) + u2d(b));
    return a > b ? a : b;
}
PGPU_DEV u64 cell_combine_dyn(int kind, u64 a, u64 b) {
    if (kind == CK_ADD_U64 || kind == CK_CNT) return a + b;
    if (kind == CK_ADD_F64) return d2u(u2d(a) + u2d(b));
    return a > b ? a : b;
}
PGPU_DEV u64 shfl_xor_u64(u64 v, int m) {
    u32 lo = __shfl_xor_sync(FULLMASK, (u32)v, m), hi = __shfl_xor_sync(FULLMASK, (u32)(v >> 32), m);
    return ((u64)hi << 32) | lo;
}
PGPU_DEV u64 shfl_u64(u64 v, int src) {
    u32 lo = __shfl_sync(FULLMASK, (u32)v, src), hi = __shfl_sync(FULLMASK, (u32)(v >> 32), src);
    return ((u64)hi << 32) | lo;
}
template <int KIND>
PGPU_DEV u64 warp_reduce(u64 v) {
    if (KIND == CK_CNT) return (u64)__reduce_add_sync(FULLMASK, (u32)v);
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) v = cell_combine<KIND>(v, shfl_xor_u64(v, m));
    return v;
}
template <int KIND>
PGPU_DEV void cell_atomic(u64* addr, u64 v) {
    if (KIND == CK_ADD_U64 || KIND == CK_CNT) { if (v) atomicAdd(addr, v); }
    else if (KIND == CK_ADD_F64) atomicAdd(reinterpret_cast<double*>(addr), u2d(v));
    else { if (v) atomicMax(addr, v); }
}

// 128-bit maximum held in two cell words [low, high] (min / max over long DECIMAL; the plan compiler maps values to an
// order-preserving unsigned image and keeps the pair 16-byte aligned)
PGPU_DEV void max128_combine(u64& alo, u64& ahi, u64 blo, u64 bhi) {
    if (bhi > ahi || (bhi == ahi && blo > alo)) {
        alo = blo;
        ahi = bhi;
    }
}
// VARCHAR(n <= 15) values travel as 16-byte cells [length byte | up to 15 bytes] in two words (lo = byte 0..7). Their order is
// the unsigned byte order of the characters, a proper prefix first (Slice.compareTo): the image puts the bytes in big-endian
// order and the length last, so that comparing the 128-bit numbers compares the strings.
PGPU_DEV u64 bswap_u64(u64 x) {
    return ((u64)__byte_perm((u32)x, 0, 0x0123) << 32) | (u64)__byte_perm((u32)(x >> 32), 0, 0x0123);
}
PGPU_DEV u128 varchar_order_image(u64 lo, u64 hi) {
    const u64 d0 = (lo >> 8) | (hi << 56);  // characters 0..7
    const u64 d1 = hi >> 8;                 // characters 8..14
    return ((u128)bswap_u64(d0) << 64) | (u128)(bswap_u64(d1) | (lo & 0xffull));
}
PGPU_DEV raw128 varchar_from_order_image(u128 img) {
    const u64 ilo = (u64)img, ihi = (u64)(img >> 64);
    const u64 d0 = bswap_u64(ihi), d1 = bswap_u64(ilo & ~0xffull);
    raw128 q;
    q.lo = (ilo & 0xffull) | (d0 << 8);
    q.hi = (d0 >> 56) | (d1 << 8);
    return q;
}
PGPU_DEV void max128_warp_reduce(u64& lo, u64& hi);   // after shfl_xor_u64
PGPU_DEV void max128_atomic(u64* cell, u64 lo, u64 hi);  // after ld_relaxed_u128 / cas_u128

// ------------------------------------------------------------------------------------------------ HBM hash table
#define TAG_EMPTY 0ull
#define TAG_LOCKED 1ull
#define TAG_FULL 2ull

PGPU_DEV u64 ld_volatile_u64(const u64* p) {
    u64 v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
PGPU_DEV u64 ld_acquire_u64(const u64* p) {
    u64 v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
PGPU_DEV void st_release_u64(u64* p, u64 v) { asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }

PGPU_DEV void st_relaxed_u64(u64* p, u64 v) { asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
PGPU_DEV u64 ld_relaxed_u64(const u64* p) {
    u64 v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// single-copy atomic 16-byte load / compare-and-swap (LDG.E.128.STRONG.GPU / ATOMG.E.CAS.128)
PGPU_DEV void ld_relaxed_u128(const u64* p, u64& lo, u64& hi) {
    asm volatile("{\n .reg .b128 v;\n ld.relaxed.gpu.global.b128 v, [%2];\n mov.b128 {%0, %1}, v;\n}" : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
}
PGPU_DEV void cas_u128(u64* p, u64 clo, u64 chi, u64 nlo, u64 nhi, u64& olo, u64& ohi) {
    asm volatile("{\n .reg .b128 c, n, o;\n mov.b128 c, {%2, %3};\n mov.b128 n, {%4, %5};\n atom.relaxed.gpu.global.cas.b128 o, [%6], c, n;\n mov.b128 {%0, %1}, o;\n}"
                 : "=l"(olo), "=l"(ohi)
                 : "l"(clo), "l"(chi), "l"(nlo), "l"(nhi), "l"(p)
                 : "memory");
}
PGPU_DEV void max128_warp_reduce(u64& lo, u64& hi) {
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) max128_combine(lo, hi, shfl_xor_u64(lo, m), shfl_xor_u64(hi, m));
}
// compare-and-swap loop: most updates do not raise the maximum and end after the load
PGPU_DEV void max128_atomic(u64* cell, u64 lo, u64 hi) {
    if ((lo | hi) == 0ull) return;  // the identity: nothing to contribute
    u64 clo, chi;
    ld_relaxed_u128(cell, clo, chi);
    while (hi > chi || (hi == chi && lo > clo)) {
        u64 olo, ohi;
        cas_u128(cell, clo, chi, lo, hi, olo, ohi);
        if (olo == clo && ohi == chi) return;
        clo = olo;
        chi = ohi;
    }
}
#define KEY_MARK 0x8000000000000000ull  // top bit of the last key word: "slot occupied" (tagless slots)

// Find the slot of `key` or claim an empty one (BigintGroupByHash.putIfAbsent :213-244, linear probing).
// Returns the slot base pointer, or nullptr when the probe chain is exhausted / the table reached its limit.
// `nins` counts the slots this thread claimed; callers add it to t.counters[0] in bulk (one atomic per warp and tile
// instead of one per group).
//   TAGLESS (keys of at most 2 words with a spare top byte): slot = [key words | cells]; the key itself is published
//   with ONE 64-/128-bit compare-and-swap, lookups are ONE relaxed load — no lock state, no fences.
//   otherwise: slot = [tag | key words | cells]; tag 0 empty, 1 locked, else 2 | hash << 8 (acquire / release protocol).
// `allow` = new slots may be claimed (false: the table reached its load-factor limit, rows needing a new slot are deferred).
// `pre` (tagless only): the home slot's key words were already loaded by probe_prefetch — a thread issues the loads of all
// its rows before it waits for the first one.
struct ProbePre {
    u64 c0, c1;
};
template <int KW, int SLOTW>
PGPU_DEV ProbePre probe_prefetch(const TableArgs& t, u64 h) {
    ProbePre p;
    const u64* s = t.slots + home_pos(h, t.capacity) * SLOTW;
    if (KW == 1) { p.c0 = ld_relaxed_u64(s); p.c1 = 0; }
    else ld_relaxed_u128(s, p.c0, p.c1);
    return p;
}
// PMASK (tagless only): bits of the last key word that take no part in the comparison — the presence byte of a bucket header.
template <int KW, int SLOTW, bool TAGLESS, u64 PMASK = 0ull>
PGPU_DEV u64* table_upsert(const TableArgs& t, const u64 (&key)[KW > 0 ? KW : 1], u64 h, bool allow, u32& nins, const ProbePre* pre = nullptr, u32 skip = 0) {
    u64 pos = home_pos(h, t.capacity) + skip;  // skip: slots of the chain the caller has already examined
    if (pos >= t.capacity) pos -= t.capacity;
    if (TAGLESS) {
        const u64 k0 = KW == 1 ? (key[0] | KEY_MARK) : key[0];
        const u64 k1 = KW == 2 ? (key[KW - 1] | KEY_MARK) : 0ull;
        bool have = pre != nullptr;
        u64 c0 = have ? pre->c0 : 0ull, c1 = have ? pre->c1 : 0ull;
#pragma unroll 1
        for (u64 probes = 0; probes < t.capacity; probes++) {
            u64* s = t.slots + pos * SLOTW;
            if (KW == 1) {
                if (!have) c0 = ld_relaxed_u64(s);
                if ((c0 & ~PMASK) == k0) return s;
                if (c0 == 0ull) {
                    if (!allow) return nullptr;
                    c0 = atomicCAS(s, 0ull, k0);
                    if (c0 == 0ull) { nins++; return s; }
                    if ((c0 & ~PMASK) == k0) return s;
                }
            } else {
                if (!have) ld_relaxed_u128(s, c0, c1);
                if (c0 == k0 && (c1 & ~PMASK) == k1) return s;
                if (c1 == 0ull) {
                    if (!allow) return nullptr;
                    cas_u128(s, 0ull, 0ull, k0, k1, c0, c1);
                    if (c1 == 0ull) { nins++; return s; }
                    if (c0 == k0 && (c1 & ~PMASK) == k1) return s;
                }
            }
            have = false;
            pos++;
            if (pos == t.capacity) pos = 0;
        }
        return nullptr;
    }
    const u64 tagv = TAG_FULL | (h << 8);
#pragma unroll 1
    for (u64 probes = 0; probes < t.capacity; probes++) {
        u64* s = t.slots + pos * SLOTW;
        for (;;) {
            u64 tag = ld_acquire_u64(s);
            if (tag == TAG_EMPTY) {
                if (!allow) return nullptr;  // full enough: defer this row
                u64 old = atomicCAS(s, TAG_EMPTY, TAG_LOCKED);
                if (old == TAG_EMPTY) {
#pragma unroll
                    for (int i = 0; i < KW; i++) s[1 + i] = key[i];
                    st_release_u64(s, tagv);
                    nins++;
                    return s;
                }
                continue;  // somebody else got it: re-read
            }
            if (tag == TAG_LOCKED) { __nanosleep(20); continue; }
            if (tag == tagv) {
                bool eq = true;
#pragma unroll
                for (int i = 0; i < KW; i++) eq &= (s[1 + i] == key[i]);
                if (eq) return s;
            }
            break;  // occupied by another key
        }
        pos++;
        if (pos == t.capacity) pos = 0;
    }
    return nullptr;
}
// slot occupied? / key words of an occupied slot (marker stripped)
template <class P>
PGPU_DEV bool slot_live(const u64* slot) {
    return P::TAGLESS ? (slot[P::KW - 1] & KEY_MARK) != 0 : (slot[0] & 0xffull) == TAG_FULL;
}
template <class P>
PGPU_DEV void slot_key(const u64* slot, u64 (&key)[P::KW > 0 ? P::KW : 1]) {
#pragma unroll
    for (int i = 0; i < P::KW; i++) key[i] = slot[P::KOFF + i];
    if (P::TAGLESS) key[P::KW - 1] &= ~KEY_MARK;
}

// ---- bucketed slots (P::BUCKETS): high-cardinality tables over ONE integer key whose values come in clusters (l_orderkey:
// 8 used keys of every 32, rows of an order adjacent). A slot holds the 8 keys that differ only in the 3 low bits of the
// key field: slot = [bucket key words | 8 x NW cells] (P::BSLOTW words). The bucket key is the packed key with those 3 bits
// cleared, published with the same single compare-and-swap as a tagless key; the last key word also carries a presence
// byte (bit P::BPRES + j: key j of the bucket exists), set with a fire-and-forget RED.OR and ignored by comparisons.
// One 160-byte fetch and write-back per 8 groups instead of eight 64 + 32-byte ones; groups are counted when the table is
// read out (body_count_groups), t.counters[0] counts claimed buckets.
// the key field (P::BKBITS bits at bit P::BSH of the packed key words) as an integer, and back
template <class P>
PGPU_DEV u64 bk_extract(const u64* w) {
    u64 k = w[0] >> P::BSH;
    if (P::BSH > 0 && P::BSH + P::BKBITS > 64) k |= w[P::KWA - 1] << (64 - P::BSH);
    if (P::BKBITS < 64) k &= (1ull << (P::BKBITS & 63)) - 1ull;
    return k;
}
template <class P>
PGPU_DEV void bk_insert(u64* w, u64 k) {
    const u64 m = P::BKBITS < 64 ? (1ull << (P::BKBITS & 63)) - 1ull : ~0ull;
    k &= m;
    w[0] = (w[0] & ~(m << P::BSH)) | (k << P::BSH);
    if (P::BSH > 0 && P::BSH + P::BKBITS > 64) w[P::KWA - 1] = (w[P::KWA - 1] & ~(m >> (64 - P::BSH))) | (k >> (64 - P::BSH));
}
template <class P>
PGPU_DEV void bucket_split(const TableArgs& t, const u64 (&key)[P::KWA], u64 (&bk)[P::KWA], u32& sub) {
#pragma unroll
    for (int i = 0; i < P::KWA; i++) bk[i] = key[i];
    const u64 k = bk_extract<P>(key) + t.bucket_off;   // wraps inside the key's width (bk_insert masks)
    sub = (u32)k & 7u;
    bk_insert<P>(bk, k & ~7ull);
}
template <class P>
PGPU_DEV void bucket_join(const TableArgs& t, const u64* slot, u32 sub, u64 (&key)[P::KWA]) {
#pragma unroll
    for (int i = 0; i < P::KW; i++) key[i] = slot[i];
    key[P::KWA - 1] &= ~(KEY_MARK | P::BPMASK);
    bk_insert<P>(key, (bk_extract<P>(key) | (u64)sub) - t.bucket_off);
}
template <class P>
PGPU_DEV u32 bucket_presence(const u64* slot) {
    const u64 w = slot[P::KWA - 1];
    return (w & KEY_MARK) ? (u32)(w >> P::BPRES) & 0xffu : 0u;
}
// update key `key` (force: claim a bucket even beyond the load-factor limit); false = deferred
template <class P>
PGPU_DEV bool bucket_update(const TableArgs& t, const u64 (&key)[P::KWA], const u64 (&hot)[P::NHA], bool force, u32& nins);

// Accumulator words come in two layouts. CELLS (P::NW words) is the persistent layout of a table slot / the global
// state; it never depends on which pages carry null vectors. HOT (P::NH words) is what the scan loop carries: cells
// whose per-row contribution is the same expression (e.g. every count when no input is NULL) share one hot word and are
// expanded back to cells only when a partial result is merged into a slot (P::atomic_cells / P::merge_cells).

// Update the group of `key` with the hot contributions `hot`. Returns false when the row has to be deferred.
PGPU_DEV bool table_may_insert(const TableArgs& t) {  // load-factor guard; armed by the host only for chunks that could reach the limit
    return t.limit == ~0ull || ld_volatile_u64(t.counters) < t.limit;
}
template <class P>
PGPU_DEV bool table_update(const TableArgs& t, const u64 (&key)[P::KW > 0 ? P::KW : 1], const u64 (&hot)[P::NHA], bool force, u32& nins) {
    u64* s = table_upsert<P::KW, P::SLOTW, P::TAGLESS>(t, key, hash_key<P::KW>(key), force || table_may_insert(t), nins);
    if (!s) return false;
    P::atomic_cells(s + P::KOFF + P::KW, hot);
    return true;
}
template <class P>
PGPU_DEV bool bucket_update(const TableArgs& t, const u64 (&key)[P::KWA], const u64 (&hot)[P::NHA], bool force, u32& nins) {
    u64 bk[P::KWA];
    u32 sub;
    bucket_split<P>(t, key, bk, sub);
    u64* s = table_upsert<P::KW, P::BSLOTW, true, P::BPMASK>(t, bk, hash_key<P::KW>(bk), force || table_may_insert(t), nins);
    if (!s) return false;
    P::atomic_cells(s + P::KW + sub * P::NW, hot);
    atomicOr(s + P::KW - 1, 1ull << (P::BPRES + sub));
    return true;
}
// add the slots a thread claimed to the table's group counter
PGPU_DEV void flush_inserts(const TableArgs& t, u32& nins) {
    if (nins) atomicAdd(t.counters, (u64)nins);
    nins = 0;
}

// Defer rows [row, row + nrows): they are re-evaluated by k_replay once the host has grown the table.
PGPU_DEV void defer_rows(const TableArgs& t, i64 row, int nrows, u32& err) {
    u64 at = atomicAdd(t.counters + 1, (u64)nrows);
    for (int k = 0; k < nrows; k++) {
        if (at + k < t.faillog_cap) t.faillog[at + k] = (u64)(row + k) + t.row_bias;
        else err |= PGPU_ERR_TABLE_FULL;
    }
}

// ------------------------------------------------------------------------------------------------ global aggregation
// One accumulator slot (no GROUP BY). Every thread reduces its rows in registers, blocks reduce through warp
// shuffles + shared memory into partials[block][NH]; the last block to finish (ticket) folds the block partials
// in block order into the persistent state, so results are deterministic run to run.
template <class P>
PGPU_DEV void body_agg_global(const PageArgs<P::NC>& a, const Scratch& sc) {
    constexpr int NH = P::NH, NHA = P::NHA;
    u64 acc[NHA];
    P::identity(acc);
    u32 err = 0;
    bool eager = false;  // warp-uniform, see body_agg_hash
    const i64 tile_rows = (i64)blockDim.x * 4;
    const i64 ntiles = (a.rows + tile_rows - 1) / tile_rows;
    for (i64 tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        i64 base = tile * tile_rows + (i64)threadIdx.x * 4;
        bool sel[4];
        u64 key[4][1];
        u64 cw[4][NHA];
        if (eager) P::eval_eager(a, base, a.rows, sel, key, cw, err);
        else P::eval(a, base, a.rows, sel, key, cw, err);
        eager = __popc(__ballot_sync(FULLMASK, sel[0] | sel[1] | sel[2] | sel[3])) >= P::EAGER_LANES;
#pragma unroll
        for (int r = 0; r < 4; r++)
            if (sel[r]) P::combine(acc, cw[r]);
    }
    __shared__ u64 sm[32][NHA];
    P::warp_reduce_cells(acc);
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    if (lane == 0) {
#pragma unroll
        for (int w = 0; w < NH; w++) sm[warp][w] = acc[w];
    }
    if (err) atomicOr(sc.err, err);
    __syncthreads();
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
        u64 tot[NHA];
        P::identity(tot);
        for (int k = 0; k < nwarps; k++) P::combine(tot, sm[k]);
#pragma unroll
        for (int w = 0; w < NH; w++) sc.partials[(u64)blockIdx.x * NHA + w] = tot[w];
        __threadfence();
        u32 t = atomicAdd(sc.ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last && threadIdx.x == 0) {
        __threadfence();
        u64 tot[NHA];
        P::identity(tot);
        for (u32 b = 0; b < gridDim.x; b++) {
            u64 part[NHA];
#pragma unroll
            for (int w = 0; w < NH; w++) part[w] = ld_volatile_u64(sc.partials + (u64)b * NHA + w);
            P::combine(tot, part);
        }
        P::merge_cells(sc.state, tot);
        *sc.ticket = 0;
    }
}

// ------------------------------------------------------------------------------------------------ hash aggregation
// Level 0 (P::RG > 0): the first RG groups a block meets live in REGISTERS: every thread keeps private accumulators
// for them and adds its rows with predicated adds — no shuffles, no shared memory, no atomics in the scan loop; the
// threads' copies are folded by warp shuffles once, when the block has consumed its last tile. TPC-H Q1 (4 groups)
// runs entirely at this level.
// Level 1 (P::USE_DIR): a block-shared key directory in shared memory maps keys to small slot numbers; each warp
// owns a private accumulator table indexed by that slot, updated without atomics after a warp-level reduction over
// lanes holding the same slot. Level 2: the HBM table, for directory misses and for the end-of-kernel flush.
template <int KW>
PGPU_DEV u32 dir_hash(const u64 (&k)[KW > 0 ? KW : 1]) {  // cheap 32-bit multiplicative hash: directory placement only
    u32 h = 0x9e3779b9u;
#pragma unroll
    for (int i = 0; i < KW; i++) h = (h ^ (u32)k[i] ^ (u32)(k[i] >> 32)) * 0x85ebca6bu;
    return h ^ (h >> 15);
}
// Returns (dense id << 16) | directory position, or -1 when the key is not in the directory and the directory is full.
// Dense ids count insertions (0, 1, 2, ... per block); ids < P::RG are the register-resident groups.
template <class P>
PGPU_DEV int dir_lookup(u64* dkeys, u32* dtags, u32* dids, u32* dpos, u32* dcount, const u64 (&key)[P::KW], u32 h) {
    constexpr int KW = P::KW, DCAP = P::DCAP;
    u32 pos = (h >> 8) & (u32)(DCAP - 1);
    const u32 tagv = 2u | (h << 8);
#pragma unroll 1
    for (int probes = 0; probes < DCAP; probes++) {
        for (;;) {
            u32 tag = *((volatile u32*)&dtags[pos]);
            if (tag == tagv) {
                bool eq = true;
#pragma unroll
                for (int i = 0; i < KW; i++) eq &= (((volatile u64*)dkeys)[pos * KW + i] == key[i]);
                if (eq) return (int)((*((volatile u32*)&dids[pos]) << 16) | pos);
                break;
            }
            if (tag == 0u) {
                if (*((volatile u32*)dcount) >= (u32)P::DLIMIT) return -1;  // directory full: go to HBM
                u32 old = atomicCAS(&dtags[pos], 0u, 1u);
                if (old == 0u) {
                    u32 id = atomicAdd(dcount, 1u);
#pragma unroll
                    for (int i = 0; i < KW; i++) dkeys[pos * KW + i] = key[i];
                    dids[pos] = id;
                    if (id < (u32)DCAP) dpos[id] = pos;
                    __threadfence_block();
                    *((volatile u32*)&dtags[pos]) = tagv;
                    return (int)((id << 16) | pos);
                }
                continue;
            }
            if (tag == 1u) continue;
            break;  // occupied by another key
        }
        pos = (pos + 1) & (u32)(DCAP - 1);
    }
    return -1;
}

template <class P>
PGPU_DEV void reg_accumulate(u64 (&racc)[P::RG > 0 ? P::RG : 1][P::NHA], int id, const u64 (&b)[P::NHA]) {
#pragma unroll
    for (int g = 0; g < P::RG; g++)
        if (id == g) P::combine(racc[g], b);
}

// Everything a thread needs to consume evaluated rows: directory + warp-private table pointers, the HBM table, the
// register-resident group accumulators. One instance per thread, fully scalarised by the compiler.
template <class P>
struct HashCtx {
    static constexpr int KW = P::KW, NH = P::NH, NHA = P::NHA, DCAP = P::DCAP, RG = P::RG;
    u64 *wacc, *dkeys, *myacc;
    u32 *dtags, *dids, *dpos, *dcount;
    u64 racc[RG > 0 ? RG : 1][NHA];
    u64 rkey[RG > 0 ? RG : 1][KW > 0 ? KW : 1];  // keys of the register-resident groups this thread has met (rmask bit g)
    u32 rmask;
    u32 nins;      // table slots claimed since the last flush_inserts
    bool use_dir;  // warp-uniform
    u32 err;
    u64* runsm;    // P::RUNS: warp-private staging of the warp's 128 rows, see consume_runs

    // shared-memory layout: wacc[nwarps][DCAP][NHA] u64 | dkeys[DCAP][KW] u64 | dtags | dids | dpos [DCAP] u32 | dcount u32
    static constexpr int dir_bytes(int nwarps) { return P::USE_DIR ? nwarps * DCAP * NHA * 8 + DCAP * KW * 8 + DCAP * 12 + 16 : 0; }
    // P::RUNS: per warp [KW + NHA][128] u64 (one array per key / hot word) + 128 run starts (1 byte each) + pad
    static constexpr int RUNS_WARP_BYTES = (KW + NHA) * 128 * 8 + 256;

    PGPU_DEV void init(unsigned char* smem, int nwarps, int warp) {
        wacc = reinterpret_cast<u64*>(smem);
        dkeys = wacc + (P::USE_DIR ? nwarps * DCAP * NHA : 0);
        dtags = reinterpret_cast<u32*>(dkeys + (P::USE_DIR ? DCAP * KW : 0));
        dids = dtags + (P::USE_DIR ? DCAP : 0);
        dpos = dids + (P::USE_DIR ? DCAP : 0);
        dcount = dpos + (P::USE_DIR ? DCAP : 0);
        myacc = wacc + (size_t)warp * DCAP * NHA;
        runsm = reinterpret_cast<u64*>(smem + (size_t)warp * RUNS_WARP_BYTES);  // P::RUNS implies !USE_DIR: the region starts at 0
        use_dir = P::USE_DIR;
        err = 0;
        rmask = 0;
        nins = 0;
#pragma unroll
        for (int g = 0; g < (RG > 0 ? RG : 1); g++) P::identity(racc[g]);
        if (P::USE_DIR) {  // whole block participates; caller synchronises afterwards
            for (int i = threadIdx.x; i < nwarps * DCAP; i += blockDim.x) P::identity_ptr(wacc + (size_t)i * NHA);
            for (int i = threadIdx.x; i < DCAP; i += blockDim.x) dtags[i] = 0u;
            if (threadIdx.x == 0) *dcount = 0u;
        }
    }

    // Consume the R evaluated rows of this thread (global row numbers base .. base+R-1). Must be called by all 32 lanes.
    template <int R>
    PGPU_DEV void consume(const TableArgs& t, i64 base, bool (&sel)[R], u64 (&key)[R][KW], u64 (&cw)[R][NHA]) {
        const int lane = threadIdx.x & 31;
        if (P::USE_DIR && use_dir) {
            // 1. directory entry of every selected row; rows of the register-resident groups are consumed right here
            int slot[R];
            int nmiss = 0, hits = 0;
            bool slow = false;
#pragma unroll
            for (int r = 0; r < R; r++) {
                slot[r] = -1;
                if (sel[r]) {
                    // fast path: the key is one of the register-resident groups this thread already knows — compare and add,
                    // no hashing, no shared memory
                    bool hit = false;
#pragma unroll
                    for (int g = 0; g < RG; g++) {
                        bool m = (rmask >> g) & 1u;
#pragma unroll
                        for (int i = 0; i < KW; i++) m &= (key[r][i] == rkey[g][i]);
                        if (m) P::combine(racc[g], cw[r]);
                        hit |= m;
                    }
                    if (hit) {
                        hits++;
                        continue;
                    }
                    const int e = dir_lookup<P>(dkeys, dtags, dids, dpos, dcount, key[r], dir_hash<KW>(key[r]));
                    if (e < 0) {  // 2. not in the (full) directory: straight to the HBM table
                        nmiss++;
                        if (!table_update<P>(t, key[r], cw[r], false, nins)) defer_rows(t, base + r, 1, err);
                    } else if (RG > 0 && (e >> 16) < RG) {
                        const int id = e >> 16;
                        reg_accumulate<P>(racc, id, cw[r]);
#pragma unroll
                        for (int g = 0; g < RG; g++) {
                            if (id == g) {
#pragma unroll
                                for (int i = 0; i < KW; i++) rkey[g][i] = key[r][i];
                            }
                        }
                        rmask |= 1u << id;
                        hits++;
                    } else {
                        slot[r] = e & 0xffff;
                        slow = true;
                    }
                }
            }
            // 3. remaining rows: one warp reduction per distinct directory slot held by this warp's 128 rows: every lane
            //    first adds up its own rows of that slot, the elected lane then adds the warp total to the warp-private
            //    table (no atomics)
            if (__any_sync(FULLMASK, slow)) {
                int rounds = 0;
#pragma unroll 1
                for (;;) {
                    int cand = -1;
#pragma unroll
                    for (int r = R - 1; r >= 0; r--) cand = slot[r] >= 0 ? slot[r] : cand;
                    u32 have = __ballot_sync(FULLMASK, cand >= 0);
                    if (!have) break;
                    if (rounds++ >= P::MAX_WARP_GROUPS) {
                        // too many distinct groups in one warp for the reduction to pay: fire-and-forget atomics instead
#pragma unroll
                        for (int r = 0; r < R; r++) {
                            if (slot[r] >= 0) {
                                if (!table_update<P>(t, key[r], cw[r], false, nins)) defer_rows(t, base + r, 1, err);
                                nmiss++;
                            }
                        }
                        break;
                    }
                    const int leader = __ffs(have) - 1;
                    const int g = __shfl_sync(FULLMASK, cand, leader);
                    u64 v[NHA];
                    P::identity(v);
#pragma unroll
                    for (int r = 0; r < R; r++) {
                        if (slot[r] == g) {
                            P::combine(v, cw[r]);
                            slot[r] = -1;
                            hits++;
                        }
                    }
                    P::warp_reduce_cells(v);
                    if (lane == leader) P::combine_ptr(myacc + (size_t)g * NHA, v);
                }
                __syncwarp();
            }
            // adaptive bypass: once most rows miss the (full) directory, stop paying for lookups. Bypassing is always
            // correct because the directory accumulators and the HBM table are merged additively at the end.
            if (__any_sync(FULLMASK, nmiss > 0)) {
                int wh = __reduce_add_sync(FULLMASK, hits), wm = __reduce_add_sync(FULLMASK, nmiss);
                if (wm > 2 * wh) use_dir = false;
            }
        } else {
            // high cardinality: fold the thread's consecutive rows with equal keys (l_orderkey-style runs) into the
            // first row of the run, then update the HBM table; span = rows folded (to defer the whole run if needed)
            int span[R];
#pragma unroll
            for (int r = 0; r < R; r++) span[r] = 1;
#pragma unroll
            for (int r = R - 1; r >= 1; r--) {
                bool same = sel[r] && sel[r - 1];
#pragma unroll
                for (int i = 0; i < KW; i++) same &= (key[r][i] == key[r - 1][i]);
                if (same) {
                    P::combine(cw[r - 1], cw[r]);
                    span[r - 1] += span[r];
                    sel[r] = false;
                }
            }
            // a run that continues in the next lane (lanes hold consecutive rows): the head run of a lane is handed to the tail
            // run of the previous lane, so a run costs one table update however it is split over threads. A single-run
            // lane that is still receiving hands over in a later pass; chains of equal keys settle right to left.
            if (P::LANE_MERGE) {
                int tr = -1;  // index of the thread's last run
#pragma unroll
                for (int r = 0; r < R; r++)
                    if (sel[r]) tr = r;
                bool tail_edge = false;  // the last run reaches the thread's last row
#pragma unroll
                for (int r = 0; r < R; r++)
                    if (r == tr) tail_edge = (r + span[r] == R);
                const i64 pbase = (i64)shfl_u64((u64)base, (lane + 31) & 31);
                const bool contig = lane > 0 && pbase + R == base;
#pragma unroll 1
                for (int pass = 0; pass < 31; pass++) {
                    u64 tk[KW];
#pragma unroll
                    for (int i = 0; i < KW; i++) tk[i] = 0;
#pragma unroll
                    for (int r = 0; r < R; r++)
                        if (r == tr) {
#pragma unroll
                            for (int i = 0; i < KW; i++) tk[i] = key[r][i];
                        }
                    const bool prev_edge = __shfl_up_sync(FULLMASK, (int)tail_edge, 1) != 0;  // every lane takes part in the shuffle
                    bool want = contig && sel[0] && prev_edge;
#pragma unroll
                    for (int i = 0; i < KW; i++) want &= (shfl_u64(tk[i], (lane + 31) & 31) == key[0][i]);
                    const bool next_wants = (__shfl_down_sync(FULLMASK, (int)want, 1) != 0) & (lane < 31);
                    const bool give = want && !(tr == 0 && next_wants);
                    if (!__any_sync(FULLMASK, give)) break;
                    const bool recv = (__shfl_down_sync(FULLMASK, (int)give, 1) != 0) & (lane < 31);
                    u64 in[NHA];
#pragma unroll
                    for (int i = 0; i < NHA; i++) in[i] = shfl_u64(cw[0][i], (lane + 1) & 31);
                    const int in_span = __shfl_down_sync(FULLMASK, span[0], 1);
                    if (recv) {
#pragma unroll
                        for (int r = 0; r < R; r++)
                            if (r == tr) {
                                P::combine(cw[r], in);
                                span[r] += in_span;
                            }
                    }
                    if (give) {
                        sel[0] = false;
                        if (tr == 0) { tr = -1; tail_edge = false; }
                    }
                }
            }
            if (P::TAGLESS) {
                const bool allow = table_may_insert(t);
                u64 h[R];
                if (allow) {
                    // three phases so that a thread never waits for one row's memory access before it has issued the next
                    // row's: (1) home-slot loads of all rows, (2) a compare-and-swap for every row whose home slot is empty,
                    // (3) results. Known-at-home and new-at-home groups (the common cases below the load-factor limit) cost
                    // two round trips per thread however many runs it holds; collisions walk the probe chain from slot + 1.
                    ProbePre pre[R];
                    u64 o0[R], o1[R];
                    int st[R];  // 0 = slot known, 1 = compare-and-swap in flight, 2 = walk the chain
#pragma unroll
                    for (int r = 0; r < R; r++) {
                        if (sel[r]) {
                            h[r] = hash_key<KW>(key[r]);
                            pre[r] = probe_prefetch<KW, P::SLOTW>(t, h[r]);
                        }
                    }
#pragma unroll
                    for (int r = 0; r < R; r++) {
                        o0[r] = o1[r] = 0;
                        st[r] = 0;
                        if (sel[r]) {
                            const u64 k0 = KW == 1 ? (key[r][0] | KEY_MARK) : key[r][0];
                            const u64 k1 = KW == 2 ? (key[r][KW - 1] | KEY_MARK) : 0ull;
                            const bool found = KW == 1 ? pre[r].c0 == k0 : (pre[r].c0 == k0 && pre[r].c1 == k1);
                            const bool empty = KW == 1 ? pre[r].c0 == 0ull : pre[r].c1 == 0ull;
                            st[r] = found ? 0 : empty ? 1 : 2;
                            if (st[r] == 1) {
                                u64* s = t.slots + home_pos(h[r], t.capacity) * P::SLOTW;
                                if (KW == 1) o0[r] = atomicCAS(s, 0ull, k0);
                                else cas_u128(s, 0ull, 0ull, k0, k1, o0[r], o1[r]);
                            }
                        }
                    }
#pragma unroll
                    for (int r = 0; r < R; r++) {
                        if (sel[r]) {
                            u64* s = t.slots + home_pos(h[r], t.capacity) * P::SLOTW;
                            if (st[r] == 1) {
                                const u64 k0 = KW == 1 ? (key[r][0] | KEY_MARK) : key[r][0];
                                const u64 k1 = KW == 2 ? (key[r][KW - 1] | KEY_MARK) : 0ull;
                                const bool claimed = KW == 1 ? o0[r] == 0ull : o1[r] == 0ull;
                                const bool found = KW == 1 ? o0[r] == k0 : (o0[r] == k0 && o1[r] == k1);
                                if (claimed) nins++;
                                else if (!found) st[r] = 2;
                            }
                            if (st[r] == 2) s = table_upsert<KW, P::SLOTW, true>(t, key[r], h[r], true, nins, nullptr, 1);
                            if (s) P::atomic_cells(s + P::KOFF + KW, cw[r]);
                            else defer_rows(t, base + r, span[r], err);
                        }
                    }
                } else {
                    // the table is at its load-factor limit: look up only (rows that need a new slot are deferred)
#pragma unroll
                    for (int r = 0; r < R; r++) {
                        if (sel[r]) {
                            h[r] = hash_key<KW>(key[r]);
                            u64* s = table_upsert<KW, P::SLOTW, true>(t, key[r], h[r], false, nins);
                            if (s) P::atomic_cells(s + P::KOFF + KW, cw[r]);
                            else defer_rows(t, base + r, span[r], err);
                        }
                    }
                }
            } else {
#pragma unroll
                for (int r = 0; r < R; r++)
                    if (sel[r] && !table_update<P>(t, key[r], cw[r], false, nins)) defer_rows(t, base + r, span[r], err);
            }
        }
        // group counter: one atomic per warp and call instead of one per claimed slot
        if (__any_sync(FULLMASK, nins != 0)) {
            const u32 tot = __reduce_add_sync(FULLMASK, nins);
            if (lane == 0) atomicAdd(t.counters, (u64)tot);
            nins = 0;
        }
    }

    // High cardinality with run compaction (P::RUNS; tagless slots, no directory). `consume` leaves the runs where the rows
    // happen to lie: a lane holds 0..4 of them, so the unrolled probe code runs four times per lane for ~1 run, and runs that
    // cross lanes need shuffle passes. Here the warp's 128 consecutive rows (lane l holds rows 4l..4l+3 of row `wbase`
    // onwards) are staged in warp-private shared memory, run starts are found with four ballots, and run j of the warp is
    // handed to lane j: the lane adds the run's rows up out of shared memory while its home-slot load is in flight and makes
    // ONE table update. Two runs per lane are in flight when a warp holds more than 32 runs (unordered keys).
    // Unselected rows form "dead" runs (start byte | 0x80) so that every row belongs to exactly one run.
    struct Run {
        int s, e;  // rows [s, e) of the warp tile
        bool act;
        u64 k[KW > 0 ? KW : 1];
        u64 h;
        ProbePre pre;
    };
    PGPU_DEV void run_issue(const TableArgs& t, const u8* starts, int nruns, int j, Run& a) {
        a.act = j < nruns;
        a.s = a.e = 0;
        if (a.act) {
            const u32 b = starts[j];
            a.s = (int)(b & 127u);
            a.e = j + 1 < nruns ? (int)(starts[j + 1] & 127u) : 128;
            a.act = (b & 128u) == 0u;
        }
        if (a.act) {
#pragma unroll
            for (int w = 0; w < KW; w++) a.k[w] = runsm[w * 128 + a.s];
            a.h = hash_key<KW>(a.k);
            a.pre = probe_prefetch<KW, P::SLOTW>(t, a.h);
        }
    }
    PGPU_DEV void run_resolve(const TableArgs& t, i64 wbase, bool allow, const Run& a) {
        if (!a.act) return;
        u64 v[NHA];
#pragma unroll
        for (int i = 0; i < NHA; i++) v[i] = runsm[(KW + i) * 128 + a.s];
#pragma unroll 1
        for (int row = a.s + 1; row < a.e; row++) {
            u64 b[NHA];
#pragma unroll
            for (int i = 0; i < NHA; i++) b[i] = runsm[(KW + i) * 128 + row];
            P::combine(v, b);
        }
        const u64 k0 = KW == 1 ? (a.k[0] | KEY_MARK) : a.k[0];
        const u64 k1 = KW == 2 ? (a.k[KW - 1] | KEY_MARK) : 0ull;
        const bool found = KW == 1 ? a.pre.c0 == k0 : (a.pre.c0 == k0 && a.pre.c1 == k1);
        const bool empty = KW == 1 ? a.pre.c0 == 0ull : a.pre.c1 == 0ull;
        u64* s = t.slots + home_pos(a.h, t.capacity) * P::SLOTW;
        if (!found) {
            if (!empty) {
                s = table_upsert<KW, P::SLOTW, true>(t, a.k, a.h, allow, nins, nullptr, 1);  // walk the chain from home + 1
            } else if (!allow) {
                s = nullptr;  // not in the table and the table is at its load-factor limit: defer
            } else {
                u64 o0 = 0, o1 = 0;
                if (KW == 1) o0 = atomicCAS(s, 0ull, k0);
                else cas_u128(s, 0ull, 0ull, k0, k1, o0, o1);
                const bool claimed = KW == 1 ? o0 == 0ull : o1 == 0ull;
                const bool same = KW == 1 ? o0 == k0 : (o0 == k0 && o1 == k1);
                if (claimed) nins++;
                else if (!same) s = table_upsert<KW, P::SLOTW, true>(t, a.k, a.h, true, nins, nullptr, 1);
            }
        }
        if (s) P::atomic_cells(s + P::KOFF + KW, v);
        else defer_rows(t, wbase + a.s, a.e - a.s, err);
    }
    // Stage the warp's 128 rows (lane l holds rows 4l .. 4l+3) in warp-private shared memory and list the run starts; returns the
    // number of runs. Must be called by all 32 lanes.
    PGPU_DEV int stage_runs(bool (&sel)[4], u64 (&key)[4][KW], u64 (&cw)[4][NHA]) {
        const int lane = threadIdx.x & 31;
        u8* starts = reinterpret_cast<u8*>(runsm + (KW + NHA) * 128);
        // 1. stage the rows: 4 consecutive 64-bit words per lane and array (two 128-bit stores)
#pragma unroll
        for (int w = 0; w < KW; w++) {
            ulonglong2* d = reinterpret_cast<ulonglong2*>(runsm + w * 128 + 4 * lane);
            d[0] = make_ulonglong2(key[0][w], key[1][w]);
            d[1] = make_ulonglong2(key[2][w], key[3][w]);
        }
#pragma unroll
        for (int i = 0; i < NHA; i++) {
            ulonglong2* d = reinterpret_cast<ulonglong2*>(runsm + (KW + i) * 128 + 4 * lane);
            d[0] = make_ulonglong2(cw[0][i], cw[1][i]);
            d[1] = make_ulonglong2(cw[2][i], cw[3][i]);
        }
        // 2. run starts: a selected row whose predecessor is unselected or has another key starts a live run, the first
        //    unselected row after a selected one a dead run; row 0 of the warp always starts one
        bool psel = __shfl_up_sync(FULLMASK, (int)sel[3], 1) != 0;
        bool peq = true;
#pragma unroll
        for (int w = 0; w < KW; w++) peq &= (shfl_u64(key[3][w], (lane + 31) & 31) == key[0][w]);
        if (lane == 0) { psel = true; peq = false; }
        bool st[4];
        st[0] = sel[0] ? !(psel && peq) : psel;
#pragma unroll
        for (int r = 1; r < 4; r++) {
            bool eq = true;
#pragma unroll
            for (int w = 0; w < KW; w++) eq &= (key[r][w] == key[r - 1][w]);
            st[r] = sel[r] ? !(sel[r - 1] && eq) : sel[r - 1];
        }
        const u32 below = (1u << lane) - 1u;
        int rank = 0, nruns = 0;
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const u32 m = __ballot_sync(FULLMASK, st[r]);
            rank += __popc(m & below);
            nruns += __popc(m);
        }
#pragma unroll
        for (int r = 0; r < 4; r++) {
            if (st[r]) {
                starts[rank] = (u8)((4 * lane + r) | (sel[r] ? 0 : 128));
                rank++;
            }
        }
        __syncwarp();
        return nruns;
    }
    // Must be called by all 32 lanes; `base` = global row number of the lane's first row, lanes hold consecutive rows.
    // INF: runs a lane keeps in flight (their home-slot loads are issued before the first one is resolved) when the warp holds
    // more than 32 runs, i.e. for keys in no order
    template <int INF = 2>
    PGPU_DEV void consume_runs(const TableArgs& t, i64 base, bool (&sel)[4], u64 (&key)[4][KW], u64 (&cw)[4][NHA]) {
        const int lane = threadIdx.x & 31;
        const i64 wbase = base - 4 * lane;
        const int nruns = stage_runs(sel, key, cw);
        const u8* starts = reinterpret_cast<const u8*>(runsm + (KW + NHA) * 128);
        // 3. one run per lane (INF in flight when the warp holds more than 32)
        const bool allow = table_may_insert(t);
#pragma unroll 1
        for (int j0 = 0; j0 < nruns; j0 += 32 * INF) {
            Run r[INF];
#pragma unroll
            for (int q = 0; q < INF; q++)
                if (q == 0 || j0 + 32 * q < nruns) run_issue(t, starts, nruns, j0 + 32 * q + lane, r[q]);  // (warp-uniform condition)
#pragma unroll
            for (int q = 0; q < INF; q++)
                if (q == 0 || j0 + 32 * q < nruns) run_resolve(t, wbase, allow, r[q]);
        }
        if (__any_sync(FULLMASK, nins != 0)) {
            const u32 tot = __reduce_add_sync(FULLMASK, nins);
            if (lane == 0) atomicAdd(t.counters, (u64)tot);
            nins = 0;
        }
        __syncwarp();  // the next tile overwrites the staging area
    }

    // ---- run compaction over a bucketed table (P::BUCKETS). As consume_runs, but the runs of one bucket — adjacent lanes,
    // because the keys arrive clustered — share ONE probe: the first lane of the bucket (the leader) finds or claims the
    // slot, its followers take the slot pointer by shuffle, every lane then adds its run to its own cells of the slot.
    struct RunB {
        int s, e;
        bool act, head, cased;  // head: first lane of its bucket; cased: the leader's compare-and-swap is in flight (pre = its result)
        int leader;
        u64 bk[KW > 0 ? KW : 1];
        u32 sub;
        u64 h;
        ProbePre pre;
    };
    // all 32 lanes. Leaders go straight for the slot with ONE compare-and-swap (P::BCASFIRST): its result says claimed / already
    // this bucket / somebody else's, so the common cases cost one memory round trip instead of a load and then the swap — and only
    // one lane in (up to) eight issues it. While the table is at its load-factor limit (allow = false) leaders only look.
    PGPU_DEV void run_issue_b(const TableArgs& t, const u8* starts, int nruns, int j, bool allow, RunB& a) {
        const int lane = threadIdx.x & 31;
        a.act = j < nruns;
        a.s = a.e = 0;
        a.sub = 0;
        a.h = 0;
        a.cased = false;
        a.pre.c0 = a.pre.c1 = 0;
#pragma unroll
        for (int w = 0; w < KW; w++) a.bk[w] = 0;
        if (a.act) {
            const u32 b = starts[j];
            a.s = (int)(b & 127u);
            a.e = j + 1 < nruns ? (int)(starts[j + 1] & 127u) : 128;
            a.act = (b & 128u) == 0u;
        }
        if (a.act) {
            u64 k[KW > 0 ? KW : 1];
#pragma unroll
            for (int w = 0; w < KW; w++) k[w] = runsm[w * 128 + a.s];
            bucket_split<P>(t, k, a.bk, a.sub);
            a.h = hash_key<KW>(a.bk);
        }
        // leaders: the first lane of each stretch of lanes that hold runs of the same bucket
        const bool pact = __shfl_up_sync(FULLMASK, (int)a.act, 1) != 0;  // every lane takes part in the shuffles
        bool same = a.act && lane > 0 && pact;
#pragma unroll
        for (int w = 0; w < KW; w++) same &= (shfl_u64(a.bk[w], (lane + 31) & 31) == a.bk[w]);
        a.head = a.act && !same;
        const u32 heads = __ballot_sync(FULLMASK, a.head);
        a.leader = a.act ? 31 - __clz(heads & ((2u << lane) - 1u)) : lane;
        if (a.head) {
            u64* s = t.slots + home_pos(a.h, t.capacity) * P::BSLOTW;
            if (P::BCASFIRST && allow) {
                const u64 k0 = KW == 1 ? (a.bk[0] | KEY_MARK) : a.bk[0];
                const u64 k1 = KW == 2 ? (a.bk[KW - 1] | KEY_MARK) : 0ull;
                if (KW == 1) a.pre.c0 = atomicCAS(s, 0ull, k0);
                else cas_u128(s, 0ull, 0ull, k0, k1, a.pre.c0, a.pre.c1);
                a.cased = true;
            } else {
                a.pre = probe_prefetch<KW, P::BSLOTW>(t, a.h);
            }
        }
    }
    PGPU_DEV void run_resolve_b(const TableArgs& t, i64 wbase, bool allow, const RunB& a) {  // all 32 lanes
        const int lane = threadIdx.x & 31;
        u64 v[NHA];
        P::identity(v);
        if (a.act) {
#pragma unroll
            for (int i = 0; i < NHA; i++) v[i] = runsm[(KW + i) * 128 + a.s];
#pragma unroll 1
            for (int row = a.s + 1; row < a.e; row++) {
                u64 b[NHA];
#pragma unroll
                for (int i = 0; i < NHA; i++) b[i] = runsm[(KW + i) * 128 + row];
                P::combine(v, b);
            }
        }
        u64* s = nullptr;
        if (a.head) {
            const u64 k0 = KW == 1 ? (a.bk[0] | KEY_MARK) : a.bk[0];
            const u64 k1 = KW == 2 ? (a.bk[KW - 1] | KEY_MARK) : 0ull;
            const bool found = KW == 1 ? (a.pre.c0 & ~P::BPMASK) == k0 : (a.pre.c0 == k0 && (a.pre.c1 & ~P::BPMASK) == k1);
            const bool empty = KW == 1 ? a.pre.c0 == 0ull : a.pre.c1 == 0ull;
            s = t.slots + home_pos(a.h, t.capacity) * P::BSLOTW;
            if (a.cased) {
                // pre = what the slot held before the swap: nothing (claimed now), this bucket, or another one
                if (empty) nins++;
                else if (!found) s = table_upsert<KW, P::BSLOTW, true, P::BPMASK>(t, a.bk, a.h, true, nins, nullptr, 1);  // walk the chain from home + 1
            } else if (!found) {
                if (!empty) {
                    s = table_upsert<KW, P::BSLOTW, true, P::BPMASK>(t, a.bk, a.h, allow, nins, nullptr, 1);
                } else if (!allow) {
                    s = nullptr;  // a new bucket while the table is at its load-factor limit: defer
                } else {
                    u64 o0 = 0, o1 = 0;
                    if (KW == 1) o0 = atomicCAS(s, 0ull, k0);
                    else cas_u128(s, 0ull, 0ull, k0, k1, o0, o1);
                    const bool claimed = KW == 1 ? o0 == 0ull : o1 == 0ull;
                    const bool mine = KW == 1 ? (o0 & ~P::BPMASK) == k0 : (o0 == k0 && (o1 & ~P::BPMASK) == k1);
                    if (claimed) nins++;
                    else if (!mine) s = table_upsert<KW, P::BSLOTW, true, P::BPMASK>(t, a.bk, a.h, true, nins, nullptr, 1);
                }
            }
        }
        s = reinterpret_cast<u64*>(shfl_u64((u64)(size_t)s, a.leader));
        u32 bit = 0;
        if (a.act) {
            if (s) {
                P::atomic_cells(s + KW + a.sub * P::NW, v);
                bit = 1u << a.sub;
            } else {
                defer_rows(t, wbase + a.s, a.e - a.s, err);
            }
        }
        // one RED.OR per bucket: the leader sets the presence bits of all its followers
        const u32 peers = __match_any_sync(FULLMASK, a.act ? a.leader : 32 + lane);
        const u32 bits = __reduce_or_sync(peers, bit);
        if (a.head && s && bits) atomicOr(s + KW - 1, (u64)bits << P::BPRES);
    }
    PGPU_DEV void consume_runs_b(const TableArgs& t, i64 base, bool (&sel)[4], u64 (&key)[4][KW], u64 (&cw)[4][NHA]) {
        const int lane = threadIdx.x & 31;
        const i64 wbase = base - 4 * lane;
        const int nruns = stage_runs(sel, key, cw);
        const u8* starts = reinterpret_cast<const u8*>(runsm + (KW + NHA) * 128);
        const bool allow = table_may_insert(t);
#pragma unroll 1
        for (int j0 = 0; j0 < nruns; j0 += 64) {
            RunB a, b;
            run_issue_b(t, starts, nruns, j0 + lane, allow, a);
            const bool two = j0 + 32 < nruns;  // warp-uniform
            if (two) run_issue_b(t, starts, nruns, j0 + 32 + lane, allow, b);
            run_resolve_b(t, wbase, allow, a);
            if (two) run_resolve_b(t, wbase, allow, b);
        }
        if (__any_sync(FULLMASK, nins != 0)) {
            const u32 tot = __reduce_add_sync(FULLMASK, nins);
            if (lane == 0) atomicAdd(t.counters, (u64)tot);
            nins = 0;
        }
        __syncwarp();  // the next tile overwrites the staging area
    }


    // End of kernel, whole block: fold register groups -> warp tables -> HBM table. `accwarps` = warps that own a wacc table
    // and hold register accumulators (the compute warps); other warps only take part in the barriers.
    PGPU_DEV void flush(const TableArgs& t, int accwarps) {
        if (!P::USE_DIR) {
            flush_inserts(t, nins);
            return;
        }
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        __syncthreads();
        if (RG > 0) {
            const int ngroups = (int)min(*((volatile u32*)dcount), (u32)RG);
            if (warp < accwarps) {
#pragma unroll
                for (int g = 0; g < RG; g++) {
                    if (g < ngroups) {  // block-uniform
                        P::warp_reduce_cells(racc[g]);
                        if (lane == 0) P::combine_ptr(myacc + (size_t)dpos[g] * NHA, racc[g]);
                    }
                }
            }
            __syncthreads();
        }
        // fold the warp-private tables in warp order and push each live directory entry to HBM
        for (int s = threadIdx.x; s < DCAP; s += blockDim.x) {
            if (dtags[s] < 2u) continue;
            u64 tot[NHA];
            P::identity(tot);
            for (int k = 0; k < accwarps; k++) P::combine_ptr_to(tot, wacc + ((size_t)k * DCAP + s) * NHA);
            u64 key[KW];
#pragma unroll
            for (int i = 0; i < KW; i++) key[i] = dkeys[s * KW + i];
            // the host keeps gridDim * DLIMIT slots in reserve below the load-factor limit for this flush
            if (!table_update<P>(t, key, tot, true, nins)) err |= PGPU_ERR_TABLE_FULL;
        }
        flush_inserts(t, nins);
    }
};

// Direct variant: every thread loads its own rows from global memory (128-bit streaming loads). Used for high
// cardinality plans (the table atomics are the bound there), for short pages and for the ragged tail of staged pages.
template <class P>
PGPU_DEV void body_agg_hash(const PageArgs<P::NC>& a, const TableArgs& t, const Scratch& sc) {
    constexpr int KW = P::KW, NHA = P::NHA;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HashCtx<P> cx;
    cx.init(smem_raw, P::NWARP, threadIdx.x >> 5);
    if (P::USE_DIR) __syncthreads();
    const i64 tile_rows = (i64)blockDim.x * 4;
    const i64 ntiles = (a.rows + tile_rows - 1) / tile_rows;
    bool eager = false;  // warp-uniform: most lanes passed the filter in the previous tile -> issue all loads at once
    for (i64 tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        i64 base = tile * tile_rows + (i64)threadIdx.x * 4;
        bool sel[4];
        u64 key[4][KW];
        u64 cw[4][NHA];
        if (eager) P::eval_eager(a, base, a.rows, sel, key, cw, cx.err);
        else P::eval(a, base, a.rows, sel, key, cw, cx.err);
        eager = __popc(__ballot_sync(FULLMASK, sel[0] | sel[1] | sel[2] | sel[3])) >= P::EAGER_LANES;
        if constexpr (P::BUCKETS) cx.consume_runs_b(t, base, sel, key, cw);
        else if (P::RUNS) cx.consume_runs(t, base, sel, key, cw);
        else cx.template consume<4>(t, base, sel, key, cw);
    }
    cx.flush(t, P::NWARP);
    if (cx.err) atomicOr(sc.err, cx.err);
}

// ------------------------------------------------------------------------------------------------ partition sweep
// A table far larger than L2 takes every update as a random DRAM round trip (10-14 G updates/s, tools/microbench/randtable.cu)
// where an L2-resident one takes 50 G/s. For long device-resident pages the host therefore reorders the rows first:
//   pgpu_k_part_count    rows per partition; partition = umulhi(hash, K), i.e. the K equal slices of the table in slot order
//                        (home_pos is monotone in the hash, whatever the capacity, so the order survives table growth)
//   (k_partition_scan, tableutil.cu: counts -> first row of every partition)
//   pgpu_k_part_scatter  copies the rows that pass the filter into partition order: the input arrays of the used channels,
//                        same encodings, so that the reordered page goes through the ordinary page kernel. A block ranks its
//                        tile by slice, then moves one array at a time through shared memory (P::move_tile)
//   pgpu_k_page_sweep    body_agg_hash over the reordered page with (a) tiles handed out by a ticket counter, so that all
//                        blocks work within a few hundred thousand rows of each other = inside one or two table slices, and
//                        (b) every tile pulling the table lines that the rows kSweepLookahead further on will touch into L2 with
//                        sequential full-line prefetches (without them each first touch is a 32-byte random DRAM read).
// Measured in isolation (tools/microbench/partsweep.cu, 600 M rows, 150 M groups, 8.6 GB table): shuffled keys 56 -> 27 ms.
struct PartArgs {
    u32* hist;     // [K] rows per partition (part_count adds, the host zeroes)
    u64* cursor;   // [K] next free row of every partition in the reordered arrays (part_scatter advances)
    u32 K;
    u32 pad;
};
struct SweepArgs {
    const u64* start;  // [K + 1] first row of every partition in the reordered batch
    u32* ticket;       // tile counter of this launch (zeroed by the host)
    i64 row0;          // batch row of this launch's row 0 (a batch is launched in chunks)
    u32 K;
    u32 pad;
};
constexpr int kPartMaxK = 1024;
constexpr int kPartTile = 4096;     // rows per scatter tile (the copy phase re-reads the tile's rows: they must still be in L1)
constexpr int kPartThreads = 256;   // => 4 quads per thread
constexpr i64 kSweepLookahead = 600000;

template <class P>
PGPU_DEV u32 part_of(const u64 (&key)[P::KW > 0 ? P::KW : 1], u32 K) {
    return (u32)__umul64hi(hash_key<P::KW>(key), (u64)K);
}

template <class P>
PGPU_DEV void body_part_count(const PageArgs<P::NC>& a, const PartArgs& pa, const Scratch& sc) {
    __shared__ u32 h[kPartMaxK];
    for (int j = threadIdx.x; j < (int)pa.K; j += blockDim.x) h[j] = 0;
    __syncthreads();
    u32 err = 0;
    const i64 tile_rows = (i64)blockDim.x * 4;
    const i64 ntiles = (a.rows + tile_rows - 1) / tile_rows;
    for (i64 tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const i64 base = tile * tile_rows + (i64)threadIdx.x * 4;
        bool sel[4];
        u64 key[4][P::KW > 0 ? P::KW : 1];
        u64 cw[4][P::NHA];
        P::eval(a, base, a.rows, sel, key, cw, err);
#pragma unroll
        for (int r = 0; r < 4; r++)
            if (sel[r]) atomicAdd(&h[part_of<P>(key[r], pa.K)], 1u);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < (int)pa.K; j += blockDim.x)
        if (h[j]) atomicAdd(&pa.hist[j], h[j]);
    if (err) atomicOr(sc.err, err);
}

// dynamic shared memory: u64 gbase[K], u32 lstart[K + 2], u32 fill[K], u16 sidx[kPartTile], u16 spid[kPartTile], then the stage of
// P::move_tile = 16 K + 8 + 4 kPartTile + kPartTile x (widest moved element) bytes (the host computes the same figure)
template <class P>
PGPU_DEV void body_part_scatter(const PageArgs<P::NC>& a, const PageArgs<P::NC>& out, const PartArgs& pa) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int K = (int)pa.K;
    u64* gbase = reinterpret_cast<u64*>(smem_raw);                 // first: keeps the 8-byte alignment
    u32* lstart = reinterpret_cast<u32*>(gbase + K);               // K + 1 (+ 1 pad)
    u32* fill = lstart + K + 2;
    unsigned short* sidx = reinterpret_cast<unsigned short*>(fill + K);
    unsigned short* spid = sidx + kPartTile;
    unsigned char* stage = reinterpret_cast<unsigned char*>(spid + kPartTile);  // kPartTile elements of the widest moved array (16-byte aligned: K is even)
    constexpr int Q = kPartTile / (kPartThreads * 4);              // quads per thread
    const int lane = threadIdx.x & 31;
    const i64 ntiles = (a.rows + kPartTile - 1) / kPartTile;
    for (i64 tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const i64 tbase = tile * kPartTile;
        for (int j = threadIdx.x; j <= K; j += kPartThreads) lstart[j] = 0;
        __syncthreads();
        unsigned short pid[Q][4];
#pragma unroll
        for (int q = 0; q < Q; q++) {
            const i64 base = tbase + ((i64)q * kPartThreads + threadIdx.x) * 4;
            bool sel[4];
            u64 key[4][P::KW > 0 ? P::KW : 1];
            u64 cw[4][P::NHA];
            u32 ignored = 0;  // the count pass has reported arithmetic errors
#pragma unroll
            for (int r = 0; r < 4; r++) sel[r] = false;
            if (base < a.rows) P::eval(a, base, a.rows, sel, key, cw, ignored);
#pragma unroll
            for (int r = 0; r < 4; r++) {
                pid[q][r] = sel[r] ? (unsigned short)part_of<P>(key[r], pa.K) : (unsigned short)0xffff;
                if (sel[r]) atomicAdd(&lstart[pid[q][r]], 1u);
            }
        }
        __syncthreads();
        if (threadIdx.x < 32) {  // exclusive scan of the K counts by one warp
            u32 carry = 0;
            for (int j0 = 0; j0 < K; j0 += 32) {
                const u32 v = j0 + lane < K ? lstart[j0 + lane] : 0u;
                u32 x = v;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const u32 y = __shfl_up_sync(FULLMASK, x, d);
                    if (lane >= d) x += y;
                }
                if (j0 + lane < K) {
                    lstart[j0 + lane] = carry + x - v;
                    fill[j0 + lane] = carry + x - v;
                }
                carry += __shfl_sync(FULLMASK, x, 31);
            }
            if (lane == 0) lstart[K] = carry;
        }
        __syncthreads();
        for (int j = threadIdx.x; j < K; j += kPartThreads) {
            const u32 tot = lstart[j + 1] - lstart[j];
            gbase[j] = tot ? atomicAdd(reinterpret_cast<unsigned long long*>(&pa.cursor[j]), (unsigned long long)tot) : 0ull;
        }
        // ranks: consecutive rows of a quad that go to the same partition take consecutive places (runs of equal keys survive)
#pragma unroll
        for (int q = 0; q < Q; q++) {
            u32 pos = 0;
#pragma unroll
            for (int r = 0; r < 4; r++) {
                if (pid[q][r] == 0xffff) continue;
                if (r > 0 && pid[q][r] == pid[q][r - 1]) {
                    pos++;
                } else {
                    u32 len = 1;
#pragma unroll
                    for (int r2 = r + 1; r2 < 4; r2++) {
                        if (pid[q][r2] != pid[q][r]) break;
                        len++;
                    }
                    pos = atomicAdd(&fill[pid[q][r]], len);
                }
                sidx[pos] = (unsigned short)((q * kPartThreads + threadIdx.x) * 4 + r);
                spid[pos] = pid[q][r];
            }
        }
        __syncthreads();
        const u32 total = lstart[K];
        const i64 left = a.rows - tbase;
        P::move_tile(a, out, tbase, (int)(left < kPartTile ? left : kPartTile), total, sidx, spid, gbase, lstart, stage);  // ends with a barrier
    }
}

#ifndef PGPU_SWEEP_WT
#define PGPU_SWEEP_WT 1024
#endif
#ifndef PGPU_SWEEP_INFLIGHT
#define PGPU_SWEEP_INFLIGHT 2
#endif
template <class P>
PGPU_DEV void body_agg_hash_sweep(const PageArgs<P::NC>& a, const TableArgs& t, const Scratch& sc, const SweepArgs& sw) {
    constexpr int KW = P::KW, NHA = P::NHA;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HashCtx<P> cx;
    cx.init(smem_raw, P::NWARP, threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    constexpr i64 WT = PGPU_SWEEP_WT;  // rows per ticket: tickets go to WARPS (a block-wide hand-out costs two barriers per tile, 30 % of the kernel)
    const i64 ntick = (a.rows + WT - 1) / WT;
    const u64 table_bytes = t.capacity * (u64)P::SLOTW * 8ull;
    const i64 batch_rows = (i64)sw.start[sw.K];
    for (;;) {
        i64 tk = 0;
        u64 pf0 = 0, pf1 = 0;
        if (lane == 0) {
            tk = (i64)atomicAdd(sw.ticket, 1u);
            // table bytes that batch rows [x0, x1) map to: a row at fraction f of partition p sits near slot (p + f) / K * capacity
            u64 b[2];
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const i64 x = sw.row0 + (tk + e) * WT + kSweepLookahead;
                if (tk >= ntick || x >= batch_rows) { b[e] = table_bytes; continue; }
                int lo = 0, hi = (int)sw.K;  // last partition that starts at or before x
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if ((i64)sw.start[mid] <= x) lo = mid; else hi = mid;
                }
                const i64 s0 = (i64)sw.start[lo], s1 = (i64)sw.start[lo + 1];
                const double f = ((double)lo + (double)(x - s0) / (double)(s1 > s0 ? s1 - s0 : 1)) / (double)sw.K;
                b[e] = (u64)(f * (double)table_bytes);
            }
            pf0 = b[0] & ~127ull;
            pf1 = b[1] > table_bytes ? table_bytes : b[1];
        }
        tk = (i64)shfl_u64((u64)tk, 0);
        if (tk >= ntick) break;
        pf0 = shfl_u64(pf0, 0);
        pf1 = shfl_u64(pf1, 0);
        for (u64 off = pf0 + (u64)lane * 128ull; off < pf1; off += 32ull * 128ull)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(t.slots) + off));
#pragma unroll 1
        for (int sub = 0; sub < (int)(WT / 128); sub++) {
            const i64 base = tk * WT + sub * 128 + lane * 4;
            bool sel[4];
            u64 key[4][KW];
            u64 cw[4][NHA];
            P::eval_eager(a, base, a.rows, sel, key, cw, cx.err);
            if (P::RUNS) cx.template consume_runs<PGPU_SWEEP_INFLIGHT>(t, base, sel, key, cw);  // (HashCtx::consume measured no faster here, even for keys in no order)
            else cx.template consume<4>(t, base, sel, key, cw);
        }
    }
    cx.flush(t, P::NWARP);
    if (cx.err) atomicOr(sc.err, cx.err);
}

// ------------------------------------------------------------------------------------------------ streaming runs
// Keys that arrive in key order (GROUP BY l_orderkey over lineitem as the table lies) need no hash table: a group is a run of
// adjacent rows. pgpu_k_runs_count + pgpu_k_runs_write put one slot-format row [key | cells] per run into a linear buffer IN ROW
// ORDER (a warp's 128 rows at the offset the scanned counts give it; a run that crosses a warp boundary leaves adjacent rows with the
// same key); pgpu_k_runs_emit merges such neighbours, finalises every group straight into the output buffers and counts the places
// where the key order DEcreases. A decrease means a key may reappear later: the host then merges the buffered rows into the hash
// table (pgpu_k_runs_to_table) and takes the ordinary path — the buffer holds partial states, nothing is lost. Any total order
// serves; P::key_order is the one a table sorted by the grouping fields comes in (first field first, SQL type order, NULLs last).
struct RunArgs {
    u64* buf;           // slot-format rows, 32-byte aligned
    u64 capacity_rows;
    u64 base_rows;      // rows the buffer already holds (earlier launches)
};

PGPU_DEV u64 shfl_up_u64(u64 v, int d) {
    return ((u64)__shfl_up_sync(FULLMASK, (u32)(v >> 32), d) << 32) | (u64)__shfl_up_sync(FULLMASK, (u32)v, d);
}

#ifndef PGPU_STREAM_SUB
#define PGPU_STREAM_SUB 1
#endif
// Two passes, no ordering between warps. One thread = R = 4 * PGPU_STREAM_SUB consecutive rows, a warp's unit = 32 R consecutive rows.
// pgpu_k_runs_count evaluates the rows and stores how many live runs end in each unit; the host side scans the counts
// (tableutil.cu runsScan) and sizes the buffer exactly; pgpu_k_runs_write evaluates the rows again and writes every unit's rows at
// its offset. A run is a maximal stretch of rows with the same key and the same filter outcome (rows the filter rejects form
// "dead" runs that are not written; keys are compared among selected rows only). Runs are cut at unit boundaries (lane 0's first
// row always starts one). Inside a thread the rows fold forwards, so that a run's total stands at its LAST row; across the lanes
// a segmented inclusive scan carries the partial total of an open run to the lane where it ends. Everything stays in registers.
// (A one-pass form, tiles ordered by tickets and a decoupled look-back, read the rows once but spent half its time in the
// look-back: 7.3 ms against 4.3 ms for 600 M rows, DESIGN.md "streaming runs".)
template <class P>
struct RunsRows {
    static constexpr int KW = P::KW, NHA = P::NHA, SUB = PGPU_STREAM_SUB, R = 4 * SUB;
    bool sel[SUB][4];
    u64 key[SUB][4][KW];
    u64 cw[SUB][4][NHA];
    u32 head, tail;

    PGPU_DEV void load(const PageArgs<P::NC>& a, i64 base, u32& err) {
#pragma unroll
        for (int q = 0; q < SUB; q++) P::eval_eager(a, base + 4 * q, a.rows, sel[q], key[q], cw[q], err);
    }
    // head bit r: row r starts a run; tail bit r: a LIVE run ends at row r
    PGPU_DEV void find_runs(int lane) {
        const bool psel = __shfl_up_sync(FULLMASK, sel[SUB - 1][3] ? 1 : 0, 1) != 0;
        bool differs = false;
#pragma unroll
        for (int w = 0; w < KW; w++) differs |= shfl_up_u64(key[SUB - 1][3][w], 1) != key[0][0][w];
        head = (lane == 0 || psel != sel[0][0] || (sel[0][0] && differs)) ? 1u : 0u;
#pragma unroll
        for (int r = 1; r < R; r++) {
            bool d = false;
#pragma unroll
            for (int w = 0; w < KW; w++) d |= key[(r - 1) >> 2][(r - 1) & 3][w] != key[r >> 2][r & 3][w];
            head |= (sel[(r - 1) >> 2][(r - 1) & 3] != sel[r >> 2][r & 3] || (sel[r >> 2][r & 3] && d)) ? 1u << r : 0u;
        }
        tail = head >> 1;
        if (__shfl_down_sync(FULLMASK, head & 1u, 1) != 0u || lane == 31) tail |= 1u << (R - 1);
#pragma unroll
        for (int r = 0; r < R; r++)
            if (!sel[r >> 2][r & 3]) tail &= ~(1u << r);
    }
    // after this the total of every run stands in the cells of its last row
    PGPU_DEV void fold(int lane) {
#pragma unroll
        for (int r = 0; r < R; r++)
            if (!sel[r >> 2][r & 3]) {
#pragma unroll
                for (int i = 0; i < NHA; i++) cw[r >> 2][r & 3][i] = 0ull;  // the identity of every cell kind
            }
#pragma unroll
        for (int r = 1; r < R; r++)
            if (!((head >> r) & 1u)) P::combine(cw[r >> 2][r & 3], cw[(r - 1) >> 2][(r - 1) & 3]);
        u64 val[NHA];
#pragma unroll
        for (int i = 0; i < NHA; i++) val[i] = cw[SUB - 1][3][i];
        bool flag = head != 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            u64 o[NHA];
#pragma unroll
            for (int i = 0; i < NHA; i++) o[i] = shfl_up_u64(val[i], d);
            const bool of = __shfl_up_sync(FULLMASK, flag ? 1 : 0, d) != 0;
            if (lane >= d && !flag) {
                P::combine(val, o);
                flag = of;
            }
        }
        u64 carry[NHA];
#pragma unroll
        for (int i = 0; i < NHA; i++) carry[i] = shfl_up_u64(val[i], 1);
        const int f = head ? __ffs(head) - 1 : R;
#pragma unroll
        for (int r = 0; r < R - 1; r++)
            if (f == r + 1) P::combine(cw[r >> 2][r & 3], carry);
        if (f == R) {
#pragma unroll
            for (int i = 0; i < NHA; i++) cw[SUB - 1][3][i] = val[i];
        }
    }
    // the live tails -> buffer rows pos, pos + 1, ...
    PGPU_DEV void write(const RunArgs& ra, u64 pos, u32& err) {
#pragma unroll
        for (int r = 0; r < R; r++) {
            if (!((tail >> r) & 1u)) continue;
            if (pos < ra.capacity_rows) {
                u64 row[P::SLOTW];
#pragma unroll
                for (int w = 0; w < P::SLOTW; w++) row[w] = 0ull;
#pragma unroll
                for (int w = 0; w < KW; w++) row[P::KOFF + w] = key[r >> 2][r & 3][w];
                row[P::KOFF + KW - 1] |= KEY_MARK;
                P::merge_cells(row + P::KOFF + KW, cw[r >> 2][r & 3]);
                ulonglong2* o = reinterpret_cast<ulonglong2*>(ra.buf + pos * P::SLOTW);  // rows are 32-byte aligned (SLOTW is a multiple of 4)
#pragma unroll
                for (int w = 0; w < P::SLOTW / 2; w++) o[w] = make_ulonglong2(row[2 * w], row[2 * w + 1]);
            } else {
                err |= PGPU_ERR_OUTPUT_FULL;
            }
            pos++;
        }
    }
};

template <class P>
PGPU_DEV void body_runs_count(const PageArgs<P::NC>& a, u32* counts, const Scratch& sc) {
    constexpr int R = RunsRows<P>::R;
    const int lane = threadIdx.x & 31;
    const i64 unit_rows = 32 * R, nunits = (a.rows + unit_rows - 1) / unit_rows;
    const i64 nw = ((i64)gridDim.x * blockDim.x) >> 5;
    u32 err = 0;
    for (i64 u = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5; u < nunits; u += nw) {
        RunsRows<P> t;
        t.load(a, u * unit_rows + (i64)lane * R, err);
        t.find_runs(lane);
        const u32 total = __reduce_add_sync(FULLMASK, (u32)__popc(t.tail));
        if (lane == 0) counts[u] = total;
    }
    if (err) atomicOr(sc.err, err);
}

template <class P>
PGPU_DEV void body_runs_write(const PageArgs<P::NC>& a, const u64* offsets, const RunArgs& ra, const Scratch& sc) {
    constexpr int R = RunsRows<P>::R;
    const int lane = threadIdx.x & 31;
    const i64 unit_rows = 32 * R, nunits = (a.rows + unit_rows - 1) / unit_rows;
    const i64 nw = ((i64)gridDim.x * blockDim.x) >> 5;
    u32 err = 0;
    for (i64 u = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5; u < nunits; u += nw) {
        RunsRows<P> t;
        t.load(a, u * unit_rows + (i64)lane * R, err);
        t.find_runs(lane);
        t.fold(lane);
        const u32 cnt = (u32)__popc(t.tail);
        u32 incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const u32 n = __shfl_up_sync(FULLMASK, incl, d);
            if (lane >= d) incl += n;
        }
        t.write(ra, ra.base_rows + offsets[u] + (incl - cnt), err);
    }
    if (err) atomicOr(sc.err, err);
}

// rows i - 1 and i of the buffer: same key? does the order decrease? (P::key_order: the grouping fields as SQL orders them)
template <class P>
PGPU_DEV void runs_compare(const u64* prev, const u64* cur, bool& same, bool& decreases) {
    const int o = P::key_order(prev + P::KOFF, cur + P::KOFF);
    same = o == 0;
    decreases = o > 0;
}

constexpr i64 kRunsMaxPieces = 64;  // buffer rows of one key the emit kernel merges serially (a piece = the key's rows inside one warp's 128)
constexpr int kRunsEmitPer = 4;     // buffer rows per thread and block chunk

// Two passes over a block's chunk of blockDim.x * kRunsEmitPer buffer rows, thread t taking rows t, t + blockDim.x, ... (a warp
// reads 32 neighbouring rows at a time). Pass 1 reads the keys (every load issued before the first use): a row is a HEAD when its
// key differs from the row before (that key comes from the lane below, lane 0 reads it), and the head's run CONTINUES when the
// row after it is no head. One atomic per chunk reserves the output rows. Pass 2 reads the heads' rows again (L2), merges the
// continuation rows and finalises.
template <class P>
PGPU_DEV void body_runs_emit(const u64* buf, i64 nrows, const OutArgs<P::NO>& out, const Scratch& sc, u32* violations) {
    constexpr int PER = kRunsEmitPer, KW = P::KW;
    __shared__ u32 s_warp[32];
    __shared__ u64 s_base;
    u32 err = 0, bad = 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    const i64 chunk = (i64)blockDim.x * PER;
    for (i64 c0 = (i64)blockIdx.x * chunk; c0 < nrows; c0 += (i64)gridDim.x * chunk) {
        u64 cur[PER][KW], before[PER][KW], after[PER][KW];
#pragma unroll
        for (int k = 0; k < PER; k++) {
            const i64 i = c0 + (i64)k * blockDim.x + threadIdx.x;
#pragma unroll
            for (int w = 0; w < KW; w++) {
                cur[k][w] = i < nrows ? buf[(u64)i * P::SLOTW + P::KOFF + w] : 0ull;
                before[k][w] = (lane == 0 && i > 0 && i < nrows) ? buf[(u64)(i - 1) * P::SLOTW + P::KOFF + w] : 0ull;
                after[k][w] = (lane == 31 && i + 1 < nrows) ? buf[(u64)(i + 1) * P::SLOTW + P::KOFF + w] : 0ull;
            }
        }
        u32 headmask = 0, contmask = 0;
#pragma unroll
        for (int k = 0; k < PER; k++) {
            const i64 i = c0 + (i64)k * blockDim.x + threadIdx.x;
            const bool valid = i < nrows;
            u64 prev[KW];
#pragma unroll
            for (int w = 0; w < KW; w++) {
                prev[w] = shfl_up_u64(cur[k][w], 1);
                if (lane == 0) prev[w] = before[k][w];
            }
            const int order = P::key_order(prev, cur[k]);
            const bool same = order == 0, dec = order > 0;
            const bool head = valid && (i == 0 || !same);
            bad += (valid && i > 0 && dec) ? 1u : 0u;
            // the row after this one: the next lane's (lane 31: the next warp's first row, read above)
            bool next_follows = __shfl_down_sync(FULLMASK, (valid && !head) ? 1 : 0, 1) != 0;
            if (lane == 31) {
                next_follows = i + 1 < nrows;
#pragma unroll
                for (int w = 0; w < KW; w++) next_follows &= after[k][w] == cur[k][w];
            }
            headmask |= head ? 1u << k : 0u;
            contmask |= (head && next_follows) ? 1u << k : 0u;
        }
        const u32 cnt = (u32)__popc(headmask);
        u32 incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const u32 n = __shfl_up_sync(FULLMASK, incl, d);
            if (lane >= d) incl += n;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const u32 w = lane < nwarps ? s_warp[lane] : 0u;
            u32 wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const u32 n = __shfl_up_sync(FULLMASK, wi, d);
                if (lane >= d) wi += n;
            }
            s_warp[lane] = wi - w;
            if (lane == 31) s_base = wi ? atomicAdd(sc.out_count, (u64)wi) : 0ull;
        }
        __syncthreads();
        i64 pos = (i64)(s_base + s_warp[warp] + incl - cnt);
#pragma unroll
        for (int k = 0; k < PER; k++) {
            if (!((headmask >> k) & 1u)) continue;
            const i64 i = c0 + (i64)k * blockDim.x + threadIdx.x;
            const u64* row = buf + (u64)i * P::SLOTW;
            u64 r[P::SLOTW];
            const ulonglong2* rv = reinterpret_cast<const ulonglong2*>(row);
#pragma unroll
            for (int w = 0; w < P::SLOTW / 2; w++) {
                const ulonglong2 q = rv[w];
                r[2 * w] = q.x;
                r[2 * w + 1] = q.y;
            }
            u64* c = r + P::KOFF + KW;
            if ((contmask >> k) & 1u) {
                for (i64 j = i + 1; j < nrows; j++) {  // the pieces of a run that was cut (warp boundaries, dead rows) follow it directly
                    const u64* nx = buf + (u64)j * P::SLOTW;
                    bool same, dec;
                    runs_compare<P>(row, nx, same, dec);
                    if (!same) break;
                    if (j - i > kRunsMaxPieces) {  // a group of thousands of rows: one thread must not walk it. The table path takes over
                        bad++;
                        break;
                    }
                    P::combine_cells(c, nx + P::KOFF + KW);
                }
            }
            if (pos < out.capacity) {
                u64 key[KW > 0 ? KW : 1];
#pragma unroll
                for (int w = 0; w < KW; w++) key[w] = r[P::KOFF + w];
                if (P::TAGLESS) key[KW - 1] &= ~KEY_MARK;
                P::emit(key, c, out, pos, err);
            } else {
                err |= PGPU_ERR_OUTPUT_FULL;
            }
            pos++;
        }
        __syncthreads();
    }
    bad = __reduce_add_sync(FULLMASK, bad);
    if (lane == 0 && bad) atomicAdd(violations, bad);
    if (err) atomicOr(sc.err, err);
}

// the buffered rows -> the hash table (the keys turned out not to be in order): rows of one key merge
template <class P>
PGPU_DEV void body_runs_to_table(const u64* buf, u64 nrows, const TableArgs& to, const Scratch& sc) {
    u32 err = 0, nins = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < nrows; i += (u64)gridDim.x * blockDim.x) {
        const u64* row = buf + i * P::SLOTW;
        u64 key[P::KW > 0 ? P::KW : 1];
        slot_key<P>(row, key);
        u64* d = table_upsert<P::KW, P::SLOTW, P::TAGLESS>(to, key, hash_key<P::KW>(key), true, nins);
        if (!d) { err |= PGPU_ERR_TABLE_FULL; continue; }
        P::atomic_cells_from(d + P::KOFF + P::KW, row + P::KOFF + P::KW);
    }
    flush_inserts(to, nins);
    if (err) atomicOr(sc.err, err);
}

// ------------------------------------------------------------------------------------------------ staged variant
// Bulk-async (TMA engine, cp.async.bulk) staging of column tiles into a shared-memory ring, mbarrier full/empty
// hand-shake, one producer warp + 2 groups of 8 consumer warps per SM. Memory-level parallelism no longer depends on
// how many loads the consumer threads can keep in flight in registers: S stages x TILE rows x bytes/row are in flight
// per SM (Q1: 4 x 46 KB) while the consumers evaluate rows out of shared memory.
PGPU_DEV u32 smem_addr(const void* p) { return (u32)__cvta_generic_to_shared(p); }
PGPU_DEV void mbar_init(u64* bar, u32 count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory"); }
PGPU_DEV void mbar_arrive(u64* bar) { asm volatile("{ .reg .b64 st; mbarrier.arrive.shared::cta.b64 st, [%0]; }" ::"r"(smem_addr(bar)) : "memory"); }
PGPU_DEV void mbar_arrive_expect_tx(u64* bar, u32 bytes) {
    asm volatile("{ .reg .b64 st; mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1; }" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
PGPU_DEV void mbar_wait(u64* bar, u32 parity) {
    u32 done = 0;
    const u32 addr = smem_addr(bar);
    while (!done) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    }
}
// global -> shared bulk copy; dst/src 16-byte aligned, bytes a multiple of 16; completes `bytes` of tx on `bar`
PGPU_DEV void bulk_g2s(void* dst, const void* src, u32 bytes, u64* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)), "l"(src), "r"(bytes),
                 "r"(smem_addr(bar))
                 : "memory");
}

// Shared-memory tile loads: the thread's R consecutive elements of a staged column (trow = first row within the tile),
// fetched with the widest aligned loads (R * sizeof(T) is a power of two; regions are 128-byte aligned).
template <typename T, int R>
PGPU_DEV void lds_tile(const unsigned char* col, int trow, T (&v)[R]) {
    constexpr int BYTES = R * (int)sizeof(T);
    const unsigned char* p = col + (size_t)trow * sizeof(T);
    union { T t[R]; int4 q[(BYTES + 15) / 16]; int2 d; u32 w; unsigned short h; u8 b; } u;
    if (BYTES >= 16) {
#pragma unroll
        for (int k = 0; k < BYTES / 16; k++) u.q[k] = *reinterpret_cast<const int4*>(p + 16 * k);
    } else if (BYTES == 8) {
        u.d = *reinterpret_cast<const int2*>(p);
    } else if (BYTES == 4) {
        u.w = *reinterpret_cast<const u32*>(p);
    } else if (BYTES == 2) {
        u.h = *reinterpret_cast<const unsigned short*>(p);
    } else {
        u.b = *p;
    }
#pragma unroll
    for (int r = 0; r < R; r++) v[r] = u.t[r];
}
// Short varchar value straight out of a staged byte region: N (<= 4) predicated byte loads, no window arithmetic
template <int N>
PGPU_DEV void lds_varchar(const unsigned char* bytes, i32 off, i32 len, u64& x, u32& err) {
    if (len > N) { err |= PGPU_ERR_VARCHAR_LEN; len = N; }
    if (len < 0) len = 0;
    u64 d = 0;
#pragma unroll
    for (int k = 0; k < N; k++)
        if (k < len) d |= (u64)bytes[off + k] << (8 * k);
    x = (u64)(u32)len | (d << 8);
}
// load_byte_window over a staged byte region (`start` is relative to the 16-byte aligned region base)
template <int MAXB>
PGPU_DEV void lds_byte_window(const unsigned char* bytes, i32 start, i32 nbytes, u64& lo, u64& hi, u32& err) {
    lo = 0; hi = 0;
    if (nbytes <= 0) return;
    if (nbytes > MAXB) { err |= PGPU_ERR_VARCHAR_LEN; nbytes = MAXB; }
    const u32* wp = reinterpret_cast<const u32*>(bytes + (start & ~3));
    const u32 sh = (u32)(start & 3) * 8u;
    const int nwords = ((start & 3) + nbytes + 3) >> 2;
    constexpr int MAXW = (MAXB + 6) / 4;
    u32 w[5] = {0u, 0u, 0u, 0u, 0u};
#pragma unroll
    for (int k = 0; k < MAXW; k++) w[k] = k < nwords ? wp[k] : 0u;
    u32 o0 = __funnelshift_r(w[0], w[1], sh);
    u32 o1 = MAXB > 4 ? __funnelshift_r(w[1], w[2], sh) : 0u;
    lo = ((u64)o1 << 32) | o0;
    if (MAXB > 8) {
        u32 o2 = __funnelshift_r(w[2], w[3], sh), o3 = __funnelshift_r(w[3], w[4], sh);
        hi = ((u64)o3 << 32) | o2;
    }
}

// P supplies: SR (rows per consumer thread), CWARPS (consumer warps), TILE = CWARPS * 32 * SR rows per stage,
// STAGE_BYTES, stage_issue(a, row0, stage, bar, lane, err) executed by the producer warp and
// eval_staged(a, stage, trow, sel, key, cw, err) executed by the consumer threads.
// Block = CWARPS consumer warps + 1 producer warp (512 threads, one block per SM, 128 registers per thread: the
// accumulators of the register-resident groups must not spill — with most of the SM's L1 carved out as shared memory
// a spill goes to L2). Every consumer warp takes its 32 * SR-row slice of every tile of the block.
// `nstages` (2..4) is chosen by the host from the shared-memory budget.
template <class P>
PGPU_DEV void body_agg_hash_staged(const PageArgs<P::NC>& a, const TableArgs& t, const Scratch& sc, int nstages) {
    constexpr int KW = P::KW, NHA = P::NHA, CW = P::CWARPS, TILE = P::TILE, SR = P::SR;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // layout: [directory + warp tables][full[4] empty[4] mbarriers][pad to 128][decoded dictionaries][stages]
    constexpr int DIRB = (HashCtx<P>::dir_bytes(CW) + 127) / 128 * 128;
    u64* full = reinterpret_cast<u64*>(smem_raw + DIRB);
    u64* empty = full + 4;
    unsigned char* dcache = smem_raw + DIRB + 128;
    unsigned char* stages = dcache + P::DCACHE_BYTES;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    HashCtx<P> cx;
    cx.init(smem_raw, CW, warp < CW ? warp : 0);
    if (threadIdx.x == 0) {
        for (int s = 0; s < 4; s++) {
            mbar_init(full + s, 1);    // the producer's arrive.expect_tx; the copies complete the transaction bytes
            mbar_init(empty + s, CW);  // one arrival per consumer warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    P::dcache_fill(a, dcache);
    __syncthreads();
    // a.rows is a whole number of tiles (the host runs the ragged tail through the direct kernel); var-width columns read
    // up to 3 offsets past the last tile, which the host keeps in range
    const i64 nfull = a.rows / TILE;
    const i64 mine = blockIdx.x < nfull ? (nfull - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;  // tiles of this block
    if (warp == CW) {
        // ---- producer
        int s = 0;
        u32 use = 0;  // how many times the ring has wrapped
        // what a tile's copies need from global memory (bounding offsets of var-width columns) is loaded two tiles ahead,
        // so the producer never sits on a load round trip between two stages
        typename P::StageLim lim;
        P::stage_limits(a, lim);
        typename P::StagePre q0, q1;
        if (mine > 0) P::stage_prefetch(a, (i64)blockIdx.x * TILE, q0);
        if (mine > 1) P::stage_prefetch(a, ((i64)blockIdx.x + gridDim.x) * TILE, q1);
        for (i64 it = 0; it < mine; it++) {
            const i64 row0 = (blockIdx.x + it * gridDim.x) * TILE;
            const typename P::StagePre cur = q0;
            q0 = q1;
            if (it + 2 < mine) P::stage_prefetch(a, (blockIdx.x + (it + 2) * gridDim.x) * TILE, q1);
            // stage_issue waits for the stage to be free, then arms the barrier and issues the copies
            P::stage_issue(a, row0, stages + (size_t)s * P::STAGE_BYTES, full + s, lane, cx.err, empty + s, use > 0, (use - 1) & 1u, cur, lim);
            if (++s == nstages) { s = 0; use++; }
        }
    } else {
        // ---- consumers
        const int trow = (warp * 32 + lane) * SR;
        int s = 0;
        u32 phase = 0;
        for (i64 it = 0; it < mine; it++) {
            mbar_wait(full + s, phase);
#pragma unroll 1
            for (int sub = 0; sub < P::SUB; sub++) {
                const int srow = sub * (TILE / P::SUB) + trow;
                const i64 base = (blockIdx.x + it * gridDim.x) * TILE + srow;
                bool sel[SR];
                u64 key[SR][KW];
                u64 cw[SR][NHA];
                P::eval_staged(a, stages + (size_t)s * P::STAGE_BYTES, dcache, srow, sel, key, cw, cx.err);
                if (sub == P::SUB - 1) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(empty + s);  // this warp is done reading the stage
                }
                cx.template consume<SR>(t, base, sel, key, cw);
            }
            if (++s == nstages) { s = 0; phase ^= 1u; }
        }
    }
    cx.flush(t, CW);
    if (cx.err) atomicOr(sc.err, cx.err);
}

// Re-run deferred rows (one row per thread) against the grown table.
template <class P>
PGPU_DEV void body_replay(const PageArgs<P::NC>& a, const TableArgs& t, const Scratch& sc, const u64* rows, u64 n) {
    u32 err = 0, nins = 0;
    for (u64 k = (u64)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (u64)gridDim.x * blockDim.x) {
        i64 row = (i64)rows[k];
        i64 base = row & ~3ll;
        bool sel[4];
        u64 key[4][P::KW > 0 ? P::KW : 1];
        u64 cw[4][P::NHA];
        u32 ignored = 0;  // arithmetic errors of this tile were already reported by the first pass
        P::eval(a, base, a.rows, sel, key, cw, ignored);
        int r = (int)(row & 3);
        if (sel[r]) {
            bool done;
            if constexpr (P::BUCKETS) done = bucket_update<P>(t, key[r], cw[r], true, nins);
            else done = table_update<P>(t, key[r], cw[r], true, nins);
            if (!done) err |= PGPU_ERR_TABLE_FULL;
        }
    }
    flush_inserts(t, nins);
    if (err) atomicOr(sc.err, err);
}

// ------------------------------------------------------------------------------------------------ ordered compaction
// Single-pass stream compaction with decoupled look-back: tiles are handed out in order by a ticket so every
// predecessor of a tile is resident or finished. tile_status[i] = (flag << 62) | value; flag 1 = tile aggregate,
// 2 = inclusive prefix. Output rows keep input order (SelectedPositions are ascending: PageFilter.java:36-48).
template <class P>
PGPU_DEV void body_compact(const PageArgs<P::NC>& a, const OutArgs<P::NO>& out, const Scratch& sc) {
    __shared__ i64 s_tile;
    __shared__ u32 s_warp[32];
    __shared__ u64 s_prefix;
    u32 err = 0;
    const i64 tile_rows = (i64)blockDim.x * 4;
    const i64 ntiles = (a.rows + tile_rows - 1) / tile_rows;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_tile = (i64)atomicAdd(sc.ticket, 1u);
        __syncthreads();
        const i64 tile = s_tile;
        if (tile >= ntiles) break;
        i64 base = tile * tile_rows + (i64)threadIdx.x * 4;
        bool sel[4];
        typename P::Proj pv[4];
        P::eval_project(a, base, a.rows, sel, pv, err);
        u32 cnt = (u32)sel[0] + (u32)sel[1] + (u32)sel[2] + (u32)sel[3];
        // block exclusive scan of cnt
        u32 incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            u32 n = __shfl_up_sync(FULLMASK, incl, d);
            if (lane >= d) incl += n;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            u32 w = lane < nwarps ? s_warp[lane] : 0u;
            u32 wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                u32 n = __shfl_up_sync(FULLMASK, wi, d);
                if (lane >= d) wi += n;
            }
            s_warp[lane] = wi - w;  // exclusive warp offsets
            u32 total = __shfl_sync(FULLMASK, wi, 31);
            // publish + look back
            u64 excl = 0;
            if (tile == 0) {
                if (lane == 0) st_release_u64(sc.tile_status, (2ull << 62) | (u64)total);
            } else {
                if (lane == 0) st_release_u64(sc.tile_status + tile, (1ull << 62) | (u64)total);
                i64 look = tile - 1;
                for (;;) {
                    i64 idx = look - lane;
                    u64 st = 0;
                    if (idx >= 0) {
                        do { st = ld_acquire_u64(sc.tile_status + idx); } while ((st >> 62) == 0);
                    } else {
                        st = 2ull << 62;  // before the first tile: inclusive prefix 0
                    }
                    u32 has_incl = __ballot_sync(FULLMASK, (st >> 62) == 2);
                    u64 val = st & 0x3fffffffffffffffull;
                    int first_incl = has_incl ? __ffs(has_incl) - 1 : 32;
                    if (lane > first_incl) val = 0;
#pragma unroll
                    for (int m = 16; m >= 1; m >>= 1) val += shfl_xor_u64(val, m);
                    excl += val;
                    if (has_incl) break;
                    look -= 32;
                }
                if (lane == 0) st_release_u64(sc.tile_status + tile, (2ull << 62) | (excl + total));
            }
            if (lane == 0) s_prefix = excl;
            if (tile == ntiles - 1 && lane == 0) *sc.out_count = excl + total;
        }
        __syncthreads();
        u64 pos = s_prefix + s_warp[warp] + (incl - cnt);
#pragma unroll
        for (int r = 0; r < 4; r++) {
            if (sel[r]) {
                if ((i64)pos < out.capacity) P::store_row(out, (i64)pos, pv[r]);
                else err |= PGPU_ERR_OUTPUT_FULL;
                pos++;
            }
        }
    }
    if (err) atomicOr(sc.err, err);
}

// ------------------------------------------------------------------------------------------------ table output / rehash
// ---- range-aware split of a table's groups (multi-GPU exchange, csrc/comm.inc) -----------------------------------------------
// Ranks own row ranges of the input; when the grouping key is clustered (l_orderkey) a rank's keys lie in a range that overlaps
// its neighbours' only at the boundaries. A group whose key lies in NO other rank's [lo, hi] exists on this rank alone: it needs no
// exchange and no second hash build, its final value is its partial state finalised. RangeArgs carries the other ranks' key
// ranges; mode 1 keeps the private groups, mode 2 the shared ones (NULL keys are always shared). Plans with P::RANGE_OK only
// (one integer grouping key).
struct RangeArgs {
    i64 lo[64], hi[64];  // key ranges of the OTHER ranks (lo > hi: that rank has no keys)
    int n;
    int mode;            // 1 = private groups only, 2 = shared groups only
};
template <class P>
PGPU_DEV i64 range_key_value(const u64* key) {  // the key field, sign-extended
    const u64 k = bk_extract<P>(key);
    if constexpr (P::BKBITS < 64) return ((i64)(k << (64 - P::BKBITS))) >> (64 - P::BKBITS);
    else return (i64)k;
}
template <class P>
PGPU_DEV bool range_pass(const u64* key, const RangeArgs& r) {
    bool shared = (key[0] & P::BNULLMASK) != 0ull;
    const i64 v = range_key_value<P>(key);
#pragma unroll 1
    for (int i = 0; i < r.n; i++) shared |= (v >= r.lo[i]) & (v <= r.hi[i]);
    return r.mode == 1 ? !shared : shared;
}
// min / max of the non-NULL keys in the table -> out[0] (atomicMin, initialised to INT64_MAX), out[1] (atomicMax, INT64_MIN)
template <class P>
PGPU_DEV void body_key_range(const TableArgs& t, i64* out) {
    i64 lo = 0x7fffffffffffffffll, hi = -0x7fffffffffffffffll - 1;
    for (u64 s = (u64)blockIdx.x * blockDim.x + threadIdx.x; s < t.capacity; s += (u64)gridDim.x * blockDim.x) {
        const u64* slot = t.slots + s * P::SLOTW;
        if (!slot_live<P>(slot)) continue;
        u64 key[P::KW > 0 ? P::KW : 1];
        slot_key<P>(slot, key);
        if (key[0] & P::BNULLMASK) continue;
        const i64 v = range_key_value<P>(key);
        lo = v < lo ? v : lo;
        hi = v > hi ? v : hi;
    }
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) {
        const i64 ol = (i64)shfl_xor_u64((u64)lo, m), oh = (i64)shfl_xor_u64((u64)hi, m);
        lo = ol < lo ? ol : lo;
        hi = oh > hi ? oh : hi;
    }
    if ((threadIdx.x & 31) == 0 && lo <= hi) {
        atomicMin(reinterpret_cast<long long*>(out), (long long)lo);
        atomicMax(reinterpret_cast<long long*>(out + 1), (long long)hi);
    }
}

// Shared groups when the overlap with the other ranks' ranges is narrow (row-range shards of a clustered table: the boundary
// orders): instead of scanning the whole table for them, look every key of the overlap intervals up. `iv` = disjoint intervals,
// start[i] = candidates before interval i; the NULL-key group is candidate number start[n] when probe_null is set.
struct ProbeArgs {
    i64 lo[64], hi[64], start[65];
    int n;
    int probe_null;
};
template <class P>
PGPU_DEV void body_emit_probe(const TableArgs& t, const OutArgs<P::NO>& out, const Scratch& sc, const ProbeArgs& pa) {
    u32 err = 0, nins = 0;
    const i64 total = pa.start[pa.n] + (pa.probe_null ? 1 : 0);
    for (i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (i64)gridDim.x * blockDim.x) {
        u64 key[P::KW > 0 ? P::KW : 1];
#pragma unroll
        for (int i = 0; i < P::KW; i++) key[i] = 0ull;
        if (idx == pa.start[pa.n]) {
            key[0] = P::BNULLMASK;
        } else {
            int i = 0;
            while (i + 1 < pa.n && idx >= pa.start[i + 1]) i++;
            bk_insert<P>(key, (u64)(pa.lo[i] + (idx - pa.start[i])));
        }
        const u64* s = table_upsert<P::KW, P::SLOTW, P::TAGLESS>(t, key, hash_key<P::KW>(key), false, nins);  // look-up only
        if (!s) continue;
        const i64 pos = (i64)atomicAdd(sc.out_count, 1ull);
        if (pos < out.capacity) P::emit(key, s + P::KOFF + P::KW, out, pos, err);
        else err |= PGPU_ERR_OUTPUT_FULL;
    }
    if (err) atomicOr(sc.err, err);
}

template <class P, bool RANGE>
PGPU_DEV void body_materialize_impl(const TableArgs& t, const OutArgs<P::NO>& out, const Scratch& sc, i64 first_slot, i64 nslots, const RangeArgs* rg);
template <class P>
PGPU_DEV void body_materialize(const TableArgs& t, const OutArgs<P::NO>& out, const Scratch& sc, i64 first_slot, i64 nslots) {
    body_materialize_impl<P, false>(t, out, sc, first_slot, nslots, nullptr);
}
template <class P>
PGPU_DEV void body_materialize_range(const TableArgs& t, const OutArgs<P::NO>& out, const Scratch& sc, i64 first_slot, i64 nslots, const RangeArgs& rg) {
    body_materialize_impl<P, true>(t, out, sc, first_slot, nslots, &rg);
}
template <class P, bool RANGE>
PGPU_DEV void body_materialize_impl(const TableArgs& t, const OutArgs<P::NO>& out, const Scratch& sc, i64 first_slot, i64 nslots, const RangeArgs* rg) {
    // A block takes blockDim * 4 consecutive slots per iteration, counts its live slots with one block scan and reserves
    // their output rows with ONE atomic on the output counter (a counter bumped per warp serialises on its L2 slice:
    // 214 M slots / 32 were 6.7 M same-address atomics, the whole cost of the kernel at 150 M groups).
    constexpr int SPT = 4;
    __shared__ u32 s_warp[32];
    __shared__ u64 s_base;
    u32 err = 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    i64 end = first_slot + nslots;
    if (end > (i64)t.capacity) end = (i64)t.capacity;
    const i64 per_iter = (i64)blockDim.x * SPT;
    for (i64 b0 = first_slot + (i64)blockIdx.x * per_iter; b0 < end; b0 += (i64)gridDim.x * per_iter) {
        const i64 s = b0 + (i64)threadIdx.x * SPT;
        bool live[SPT];
        u32 cnt = 0;
#pragma unroll
        for (int k = 0; k < SPT; k++) {
            live[k] = s + k < end && slot_live<P>(t.slots + (u64)(s + k) * P::SLOTW);
            if (RANGE && live[k]) {
                u64 fk[P::KW > 0 ? P::KW : 1];
                slot_key<P>(t.slots + (u64)(s + k) * P::SLOTW, fk);
                live[k] = range_pass<P>(fk, *rg);
            }
            cnt += live[k] ? 1u : 0u;
        }
        u32 incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const u32 n = __shfl_up_sync(FULLMASK, incl, d);
            if (lane >= d) incl += n;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const u32 w = lane < nwarps ? s_warp[lane] : 0u;
            u32 wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const u32 n = __shfl_up_sync(FULLMASK, wi, d);
                if (lane >= d) wi += n;
            }
            s_warp[lane] = wi - w;  // exclusive prefix of the warp totals
            if (lane == 31) s_base = wi ? atomicAdd(sc.out_count, (u64)wi) : 0ull;
        }
        __syncthreads();
        i64 pos = (i64)(s_base + s_warp[warp] + incl - cnt);
#pragma unroll
        for (int k = 0; k < SPT; k++) {
            if (live[k]) {
                if (pos < out.capacity) {
                    const u64* slot = t.slots + (u64)(s + k) * P::SLOTW;
                    u64 key[P::KW > 0 ? P::KW : 1];
                    slot_key<P>(slot, key);
                    P::emit(key, slot + P::KOFF + P::KW, out, pos, err);
                } else {
                    err |= PGPU_ERR_OUTPUT_FULL;
                }
                pos++;
            }
        }
        __syncthreads();  // s_warp / s_base are rewritten by the next iteration
    }
    if (err) atomicOr(sc.err, err);
}

template <class P>
PGPU_DEV void body_rehash(const TableArgs& from, const TableArgs& to, const Scratch& sc) {
    u32 err = 0, nins = 0;
    for (u64 s = (u64)blockIdx.x * blockDim.x + threadIdx.x; s < from.capacity; s += (u64)gridDim.x * blockDim.x) {
        const u64* slot = from.slots + s * P::SLOTW;
        if (!slot_live<P>(slot)) continue;
        u64 key[P::KW > 0 ? P::KW : 1];
        slot_key<P>(slot, key);
        u64* d = table_upsert<P::KW, P::SLOTW, P::TAGLESS>(to, key, hash_key<P::KW>(key), true, nins);
        if (!d) { err |= PGPU_ERR_TABLE_FULL; continue; }
#pragma unroll
        for (int w = 0; w < P::NW; w++) d[P::KOFF + P::KW + w] = slot[P::KOFF + P::KW + w];  // keys are distinct: plain stores
    }
    flush_inserts(to, nins);
    if (err) atomicOr(sc.err, err);
}

// ---- bucketed table: output, growth, group count -------------------------------------------------------------------------
// One thread per bucket: its live keys (presence bits) become consecutive output rows.
template <class P>
PGPU_DEV void body_materialize_b(const TableArgs& t, const OutArgs<P::NO>& out, const Scratch& sc, i64 first_slot, i64 nslots) {
    __shared__ u32 s_warp[32];
    __shared__ u64 s_base;
    u32 err = 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    i64 end = first_slot + nslots;
    if (end > (i64)t.capacity) end = (i64)t.capacity;
    for (i64 b0 = first_slot + (i64)blockIdx.x * blockDim.x; b0 < end; b0 += (i64)gridDim.x * blockDim.x) {
        const i64 s = b0 + threadIdx.x;
        const u64* slot = t.slots + (u64)s * P::BSLOTW;
        const u32 pres = s < end ? bucket_presence<P>(slot) : 0u;
        const u32 cnt = (u32)__popc(pres);
        u32 incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const u32 n = __shfl_up_sync(FULLMASK, incl, d);
            if (lane >= d) incl += n;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const u32 w = lane < nwarps ? s_warp[lane] : 0u;
            u32 wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const u32 n = __shfl_up_sync(FULLMASK, wi, d);
                if (lane >= d) wi += n;
            }
            s_warp[lane] = wi - w;
            if (lane == 31) s_base = wi ? atomicAdd(sc.out_count, (u64)wi) : 0ull;
        }
        __syncthreads();
        i64 pos = (i64)(s_base + s_warp[warp] + incl - cnt);
#pragma unroll 1
        for (u32 j = 0; j < 8; j++) {
            if (!((pres >> j) & 1u)) continue;
            if (pos < out.capacity) {
                u64 key[P::KWA];
                bucket_join<P>(t, slot, j, key);
                P::emit(key, slot + P::KW + j * P::NW, out, pos, err);
            } else {
                err |= PGPU_ERR_OUTPUT_FULL;
            }
            pos++;
        }
        __syncthreads();
    }
    if (err) atomicOr(sc.err, err);
}

template <class P>
PGPU_DEV void body_rehash_b(const TableArgs& from, const TableArgs& to, const Scratch& sc) {
    u32 err = 0, nins = 0;
    for (u64 s = (u64)blockIdx.x * blockDim.x + threadIdx.x; s < from.capacity; s += (u64)gridDim.x * blockDim.x) {
        const u64* slot = from.slots + s * P::BSLOTW;
        const u64 last = slot[P::KWA - 1];
        if (!(last & KEY_MARK)) continue;
        u64 bk[P::KWA];
#pragma unroll
        for (int i = 0; i < P::KW; i++) bk[i] = slot[i];
        bk[P::KWA - 1] &= ~(KEY_MARK | P::BPMASK);
        u64* d = table_upsert<P::KW, P::BSLOTW, true, P::BPMASK>(to, bk, hash_key<P::KW>(bk), true, nins);
        if (!d) { err |= PGPU_ERR_TABLE_FULL; continue; }
        d[P::KWA - 1] = last;  // buckets are distinct: plain stores (the presence byte travels with the header)
#pragma unroll 1
        for (int w = 0; w < 8 * P::NW; w++) d[P::KW + w] = slot[P::KW + w];
    }
    flush_inserts(to, nins);
    if (err) atomicOr(sc.err, err);
}

// groups in the table = presence bits set; added to *sc.out_count
template <class P>
PGPU_DEV void body_count_groups(const TableArgs& t, const Scratch& sc) {
    u32 n = 0;
    for (u64 s = (u64)blockIdx.x * blockDim.x + threadIdx.x; s < t.capacity; s += (u64)gridDim.x * blockDim.x)
        n += (u32)__popc(bucket_presence<P>(t.slots + s * P::BSLOTW));
    n = __reduce_add_sync(FULLMASK, n);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(sc.out_count, (u64)n);
}

// ---- gather exchange (small results): FINAL over the partial pages the peers pushed into this GPU's inbox ------------------------
// Stands in for ExchangeOperator -> HashAggregationOperator(Step.FINAL).addInput over every upstream task's partial page
// (main/operator/HashAggregationOperator.java:380-412 with Step.FINAL: addIntermediate = deserialize + combine). The host is out
// of the loop: every rank's push kernel (csrc/exchange.cu) stores its partial rows and then a header {sequence, rows, error bits}
// into the inbox slot it owns on every GPU (peer memory over NVLink, st.release.sys); this kernel — one block — waits for the
// header of each slot in turn and runs the plan's page kernel body over the slot's rows, which lie in LOCAL HBM.
PGPU_DEV u64 ld_acquire_sys_u64(const u64* p) {
    u64 v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
PGPU_DEV u64 globaltimer_ns() {
    u64 t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define PGPU_ERR_PEER_TIMEOUT 128u  // a peer's partial page did not arrive (the peer failed)
struct GatherArgs {
    const u64* hdr;        // header of slot 0: word 0 sequence, word 1 rows, word 2 sender's error bits
    int npages;
    int hdr_stride_words;  // 64-bit words from one slot's header to the next
    u64 seq;
    u64 timeout_ns;
};
// thread 0 waits for slot i's header; returns the slot's row count to every thread
PGPU_DEV i64 gather_wait(const GatherArgs& g, int i, const Scratch& sc) {
    __shared__ i64 s_rows;
    if (threadIdx.x == 0) {
        const u64* h = g.hdr + (size_t)i * (size_t)g.hdr_stride_words;
        const u64 t0 = globaltimer_ns();
        i64 rows = 0;
        u32 err = 0;
        for (;;) {
            if (ld_acquire_sys_u64(h) == g.seq) {
                rows = (i64)h[1];
                err = (u32)h[2];
                break;
            }
            if (globaltimer_ns() - t0 > g.timeout_ns) { err = PGPU_ERR_PEER_TIMEOUT; break; }
            __nanosleep(200);
        }
        if (err) atomicOr(sc.err, err);
        s_rows = err ? 0 : rows;
    }
    __syncthreads();
    const i64 r = s_rows;
    __syncthreads();
    return r;
}
template <class P>
PGPU_DEV void body_gather_hash(const PageArgs<P::NC>* pages, const GatherArgs& g, const TableArgs& t, const Scratch& sc) {
    for (int i = 0; i < g.npages; i++) {
        const i64 rows = gather_wait(g, i, sc);
        if (rows > 0) {
            PageArgs<P::NC> a = pages[i];
            a.rows = rows;
            body_agg_hash<P>(a, t, sc);
        }
        __syncthreads();
    }
}
template <class P>
PGPU_DEV void body_gather_global(const PageArgs<P::NC>* pages, const GatherArgs& g, const Scratch& sc) {
    for (int i = 0; i < g.npages; i++) {
        const i64 rows = gather_wait(g, i, sc);
        if (rows > 0) {
            PageArgs<P::NC> a = pages[i];
            a.rows = rows;
            body_agg_global<P>(a, sc);
        }
        __syncthreads();
    }
}

// Single-slot output of a global aggregation (AggregationOperator.getOutput, operator/AggregationOperator.java:164-186).
template <class P>
PGPU_DEV void body_emit_global(const Scratch& sc, const OutArgs<P::NO>& out) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        u32 err = 0;
        u64 nokey[1] = {0};
        P::emit(nokey, sc.state, out, 0, err);
        *sc.out_count = 1;
        if (err) atomicOr(sc.err, err);
    }
}
#endif  // PGPU_HOST_EMULATION
