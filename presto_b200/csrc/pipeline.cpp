// pipeline.cpp — the C ABI of include/prestogpu.h: one native operator instance per pgpu_pipeline.
//
// Host-side state machine follows the reference operators it stands in for:
//   needsInput/addInput/getOutput/finish/isFinished  main/operator/HashAggregationOperator.java:367-518,
//                                                    main/operator/AggregationOperator.java:120-186,
//                                                    main/operator/WorkProcessorOperatorAdapter.java:170-209 (filter/project)
//   table growth                                     main/operator/BigintGroupByHash.java:262-313 (tryRehash)
// Device work is queued on one CUDA stream per pipeline; nothing here is thread-local, so the Driver may call
// from a different thread each quantum (main/operator/Driver.java:284-305).
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <chrono>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/prestogpu.h"
#include "codegen.h"
#include "exchange.h"
#include "hostcopy.h"
#include "jit.h"
#include "plan.h"
#include "serde.h"
#include "tableutil.h"
#include "parquet.h"
#include "parquet_host.h"
#include "parquet_ingest.h"

using namespace pgpu;

namespace {

// device-side structs (must match device/pgpu_device.cuh)
struct ColArg {
    const void* values;
    const uint8_t* nulls;
    const int32_t* offsets;
    const int32_t* ids;
    const void* dvalues;
    const uint8_t* dnulls;
    const int32_t* doffsets;
    int32_t dcount;
    int32_t pad;
};
struct TableArgs {
    uint64_t* slots;
    uint64_t capacity;
    uint64_t* counters;
    uint64_t limit;
    uint64_t* faillog;
    uint64_t faillog_cap;
    uint64_t row_bias;
    uint64_t bucket_off;
};
struct Scratch {
    uint32_t* err;
    uint64_t* partials;
    uint32_t* ticket;
    uint64_t* state;
    uint64_t* tile_status;
    uint64_t* out_count;
};
struct OutCol {
    void* values;
    uint8_t* nulls;
};
struct RangeArgs {
    int64_t lo[64], hi[64];
    int32_t n, mode;
};
struct ProbeArgs {
    int64_t lo[64], hi[64], start[65];
    int32_t n, probe_null;
};
struct RunArgs {
    uint64_t* buf;
    uint64_t capacity_rows;
    uint64_t base_rows;
};
struct PartArgs {
    uint32_t* hist;
    uint64_t* cursor;
    uint32_t K;
    uint32_t pad;
};
struct SweepArgs {
    const uint64_t* start;
    uint32_t* ticket;
    int64_t row0;
    uint32_t K;
    uint32_t pad;
};
struct GatherArgs {
    const uint64_t* hdr;
    int32_t npages;
    int32_t hdr_stride_words;
    uint64_t seq;
    uint64_t timeout_ns;
};

constexpr int kThreads = 256;
constexpr int64_t kTileRows = kThreads * 4;
constexpr int64_t kHashChunkRows = 64ll << 20;
// partition sweep (device header, "partition sweep"): same constants as the kernels
constexpr int kPartMaxK = 1024, kPartTile = 4096, kPartThreads = 256;
constexpr size_t partScatterSmem(int K, int stageWidth) { return (size_t)16 * K + 8 + (size_t)4 * kPartTile + (size_t)kPartTile * stageWidth; }
constexpr int64_t kSweepMinRows = 16ll << 20;          // shorter batches do not pay for a pass over the table
constexpr size_t kSweepMinTableBytes = 256ull << 20;   // a table that is not much larger than L2 (126 MB) is fast as it is
constexpr size_t kSweepSliceBytes = 32ull << 20;       // table bytes per partition: what stays L2 resident while a partition is consumed  // rows per launch of the hash-aggregation kernel (bounds the deferred-row log)
constexpr int64_t kHostChunkRows = 8ll << 20;   // rows per staged H2D chunk

thread_local std::string g_lastGlobalError;

struct Variant {
    GenResult gen;
    cudaLibrary_t lib = nullptr;
    cudaKernel_t kPage = nullptr, kEmit = nullptr, kRehash = nullptr, kReplay = nullptr, kPageStaged = nullptr, kCount = nullptr, kGather = nullptr, kEmitRange = nullptr, kKeyRange = nullptr, kEmitProbe = nullptr;
    cudaKernel_t kPartCount = nullptr, kPartScatter = nullptr, kPageSweep = nullptr;  // partition sweep (gen.sweepOk)
    cudaKernel_t kRunsCount = nullptr, kRunsWrite = nullptr, kRunsEmit = nullptr, kRunsToTable = nullptr;   // streaming runs (gen.streamOk)
    int streamBlocksPerSM = 1;
    int sweepBlocksPerSM = 1, scatterBlocksPerSM = 1;
    int blocksPerSM = 1;
    int stagedStages = 0;      // stages of the shared-memory ring (0 = staged variant unavailable)
    size_t stagedSmem = 0;
};

struct DeviceBuf {
    void* ptr = nullptr;
    size_t cap = 0;
};

struct StageSlot {
    DeviceBuf dev;
    void* host = nullptr;  // pinned
    size_t hostCap = 0;
    cudaEvent_t done = nullptr;  // kernel that consumed the slot finished
    bool inFlight = false;
};

// Where the arrays of one input channel lie inside a staging slot (same byte offsets in the pinned host half and the device
// half). A slot holds ONE super-page: the rows of as many consecutive small input pages as fit, appended array by array.
struct StageCol {
    size_t values = 0, nulls = 0, offsets = 0, ids = 0, dvalues = 0, dnulls = 0, doffsets = 0;
    size_t valuesCap = 0;  // var-width: payload bytes the region holds
    int width = 0;         // fixed width: bytes per position
    int encoding = PGPU_ENC_PLAIN;
    bool nullable = false; // the block (PLAIN) / its dictionary carries a null vector
};
struct StageLayout {
    int64_t capRows = 0;
    size_t bytes = 0;
    std::vector<StageCol> cols;  // per input channel (used channels only)
};
// The super-page under construction in slots[nextSlot].
struct PendingPage {
    Variant* v = nullptr;
    std::string key;                 // page signature of the rows gathered so far
    const StageLayout* lay = nullptr;
    int64_t rows = 0;
    std::vector<int64_t> vbytes;     // per channel: var-width payload bytes so far
    std::vector<int32_t> dcount;     // per channel: entries of the dictionary / RLE value block held in the slot
    std::vector<size_t> dbytes;      // per channel: bytes of its values
};

// One channel of a host page as the appender reads it: pgpu_column arrays or the pieces of a serialized block (which may be
// unaligned and carry END offsets without the leading 0, spi/block/VariableWidthBlockEncoding.java:37-58).
struct HostCol {
    int encoding = PGPU_ENC_PLAIN;
    const void* values = nullptr;
    const uint8_t* nulls = nullptr;
    const char* offsets = nullptr;   // int32 entries
    bool endOffsets = false;
    const char* ids = nullptr;       // int32 entries
    // dictionary / RLE value block
    const void* dvalues = nullptr;
    const uint8_t* dnulls = nullptr;
    const char* doffsets = nullptr;
    bool dEndOffsets = false;
    int64_t dcount = 0;
};

constexpr size_t kDictStageBytes = 64u << 10;  // a coalesced super-page carries one dictionary per channel, up to this size
constexpr int kOffsetsSlack = 8;               // offsets entries kept valid past the end of a super-page (the staged kernel reads 4)

struct HostOut {
    void* ptr = nullptr;
    size_t cap = 0;
};

}  // namespace

struct pgpu_pipeline {
    std::unique_ptr<Plan> plan;
    int device = -1;
    int numSMs = 148;
    cudaStream_t stream = nullptr;
    std::map<std::string, Variant> variants;
    GenResult layout;  // layout fields (KW/NW/SLOTW/out/mode) are signature independent
    bool haveLayout = false;
    std::string lastSource;
    std::string err;
    pgpu_stats stats{};
    bool timing = false;
    cudaEvent_t evA = nullptr, evB = nullptr;

    // operator state
    bool finishing = false, finished = false;
    bool sawInput = false;
    bool outputReady = false;      // COMPACT: a page result is pending; aggregation: result materialised
    int64_t outRowsTotal = 0, outRowsServed = 0;
    // COMPACT with merge-min-rows / merge-min-bytes (MergePages): rows [0, mergeRows) of the output arena are buffered small
    // pages; the next page's kernel writes behind them. compactPending: a page kernel is in flight, its row count unread.
    // outBoundary: an output page ends here even if more rows follow (the flushed buffer, then the big page that flushed it)
    bool compactPending = false;
    int64_t mergeRows = 0, outBoundary = 0, outRowBytes = 0;

    // device scratch
    uint32_t* dErr = nullptr;
    uint32_t* dTicket = nullptr;
    uint64_t* dOutCount = nullptr;
    uint64_t* dCounters = nullptr;
    uint64_t* dState = nullptr;
    uint64_t* dPartials = nullptr;
    size_t partialsCap = 0;
    DeviceBuf tileStatus;
    DeviceBuf failLog;   // deferred rows of the current hash-aggregation chunk
    // streaming runs (device header, "streaming runs"): keys in key order are aggregated run by run into a linear buffer of
    // slot-format rows, no hash table. STREAM_OFF once the order was seen to break (the buffer then went into the table).
    enum StreamMode { STREAM_UNDECIDED, STREAM_ON, STREAM_OFF } streamMode = STREAM_UNDECIDED;
    DeviceBuf runBuf, runCounts, runOffsets;
    int64_t runRows = 0;          // rows the buffer holds
    uint32_t* dViolations = nullptr;
    DeviceBuf sweepBuf;  // partition sweep: the reordered arrays of one batch
    DeviceBuf sweepMeta; // ... [hist u32 x K | ticket u32 x 64 | cursor u64 x K | start u64 x (K + 1)]
    uint64_t* hostScalars = nullptr;  // pinned, 32 words: [0] out_count, [1..2] counters, [3] err, [4..12] key-clustering probe

    // table
    uint64_t* dSlots = nullptr;
    uint64_t capacity = 0;
    // slotsKnown: table slots claimed at the last look (what the load factor is about); groupsKnown: groups in the table (what the
    // output is sized by). The same number for one-key-per-slot tables; a bucketed table claims one slot per 8 neighbouring keys.
    int64_t slotsKnown = 0, groupsKnown = 0, rowsSinceKnown = 0;
    enum { TABLE_UNDECIDED = 0, TABLE_PER_KEY = 1, TABLE_BUCKETS = 2 };
    uint64_t bucketOff = 0;        // bucketed table: buckets are cut at 8n - bucketOff (chosen by the probe; PGPU_TUNE_BUCKET_OFF)
    double bucketOccupancy = 8.0;  // keys per bucket the probe expects (table sizing)
    int tableMode = TABLE_UNDECIDED;  // decided by the first page (key clustering probe) or PGPU_TUNE_BUCKETS; survives pgpu_reset

    // staging + output
    StageSlot slots[3];
    int nextSlot = 0;
    PendingPage pend;
    // multi-GPU exchange, range-aware split (comm.inc): the same plan as a SINGLE step — its emit kernel finalises groups of THIS
    // (PARTIAL-step) pipeline's table that no other rank can hold; rows the exchange appends to a FINAL pipeline's output
    pgpu_pipeline* singleTwin = nullptr;
    int64_t extraOutRows = 0;
    int64_t tableSizeHint = 0;  // groups to size a not yet allocated table for instead of the plan's expected-groups (FINAL behind a range-split exchange)
    std::map<std::string, StageLayout> stageLayouts;
    size_t stageTargetBytes = 64u << 20;  // PGPU_STAGE_MB
    // output columns live in ONE device allocation (values / nulls of column c at outOffV[c] / outOffN[c]) so that a
    // small result reaches the host with a single copy instead of two per column
    DeviceBuf outArena;
    HostOut outHostArena;
    std::vector<size_t> outOffV, outOffN;
    size_t outLayoutBytes = 0;
    std::vector<DeviceBuf> outDev;       // per output column: values (views into outArena)
    std::vector<DeviceBuf> outDevNulls;
    std::vector<HostOut> outHost, outHostNulls, outHostAux;  // aux: offsets for var-width
    int64_t outCapacityRows = 0;
    // serialized-page ingest: the payload + the decoded, aligned block arrays of the page being added
    DeviceBuf serdeArena;
    DeviceBuf serdeOutArena;            // getOutput in the wire format: packed pieces of every channel
    HostOut serdeHost, serdeTotals;     // the payload (pinned) and the per-channel counts the layout depends on
    size_t serdeLen = 0;
    DeviceBuf parquetArena;             // Parquet ingest: chunk bytes, run tables and the decoded block arrays of the row group being added
    // partitioned output (exchange)
    // one allocation for every partitioned column (a peer maps it with a single IPC handle); arenas replaced by a
    // larger one stay allocated until destroy because a peer may still have them mapped
    DeviceBuf partArena;
    std::vector<void*> partRetired;
    std::vector<void*> partDev, partDevNulls;  // per output column: pointers into partArena
    DeviceBuf partPid, partDst, partCounts;
    unsigned long long* hostPart = nullptr;  // pinned: counts[64] + cursors[64]
};

namespace {

#define CU(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) {                                                                          \
            p->err = std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #call;                 \
            return PGPU_CUDA_ERROR;                                                                       \
        }                                                                                                 \
    } while (0)

pgpu_status fail(pgpu_pipeline* p, pgpu_status st, const std::string& m) {
    p->err = m;
    return st;
}

size_t align256(size_t x) { return (x + 255) / 256 * 256; }

pgpu_status ensureDev(pgpu_pipeline* p, DeviceBuf& b, size_t bytes) {
    if (b.cap >= bytes && b.ptr) return PGPU_OK;
    if (b.ptr) {
        CU(cudaStreamSynchronize(p->stream));
        CU(cudaFree(b.ptr));
        b.ptr = nullptr;
        b.cap = 0;
    }
    size_t want = std::max<size_t>(bytes, 256);
    want = (want + 255) / 256 * 256;
    CU(cudaMalloc(&b.ptr, want));
    b.cap = want;
    return PGPU_OK;
}

pgpu_status ensureHost(pgpu_pipeline* p, HostOut& b, size_t bytes) {
    if (b.cap >= bytes && b.ptr) return PGPU_OK;
    if (b.ptr) {
        CU(cudaFreeHost(b.ptr));
        b.ptr = nullptr;
        b.cap = 0;
    }
    size_t want = std::max<size_t>(bytes, 64);
    CU(cudaMallocHost(&b.ptr, want));
    b.cap = want;
    return PGPU_OK;
}

std::string sigKey(const std::vector<ColSig>& s) {
    std::string k;
    for (auto& c : s) {
        k.push_back((char)('0' + c.encoding));
        k.push_back(c.nullable ? 'n' : '-');
    }
    return k;
}

pgpu_status getVariant(pgpu_pipeline* p, const std::vector<ColSig>& sig, Variant** out) {
    const bool buckets = p->tableMode == pgpu_pipeline::TABLE_BUCKETS;
    std::string key = sigKey(sig) + (buckets ? "B" : "");
    auto it = p->variants.find(key);
    if (it != p->variants.end()) {
        *out = &it->second;
        return PGPU_OK;
    }
    Variant v;
    try {
        GenOptions opt;
        opt.buckets = buckets;
        v.gen = generate(*p->plan, sig, opt);
        JitResult jr = jitCompile(v.gen.source);
        if (jr.cacheHit) p->stats.jit_cache_hits++;
        else p->stats.jit_compiles++;
        p->lastSource = v.gen.source;
        if (!p->haveLayout) {
            p->layout = v.gen;
            p->layout.source.clear();
            p->haveLayout = true;
        } else if (p->layout.KW != v.gen.KW || p->layout.NW != v.gen.NW || p->layout.SLOTW != v.gen.SLOTW || p->layout.cellKinds != v.gen.cellKinds) {
            return fail(p, PGPU_COMPILER_ERROR, "internal: accumulator layout depends on the page signature");
        }
        if (p->device >= 0) {
            CU(cudaSetDevice(p->device));
            CU(cudaLibraryLoadData(&v.lib, jr.cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0));
            CU(cudaLibraryGetKernel(&v.kPage, v.lib, "pgpu_k_page"));
            if (v.gen.mode != KernelMode::COMPACT) CU(cudaLibraryGetKernel(&v.kEmit, v.lib, "pgpu_k_emit"));
            if (v.gen.mode == KernelMode::AGG_HASH) {
                CU(cudaLibraryGetKernel(&v.kRehash, v.lib, "pgpu_k_rehash"));
                CU(cudaLibraryGetKernel(&v.kReplay, v.lib, "pgpu_k_replay"));
                if (v.gen.buckets) CU(cudaLibraryGetKernel(&v.kCount, v.lib, "pgpu_k_count"));
                if (v.gen.rangeOk) {
                    CU(cudaLibraryGetKernel(&v.kEmitRange, v.lib, "pgpu_k_emit_range"));
                    CU(cudaLibraryGetKernel(&v.kKeyRange, v.lib, "pgpu_k_keyrange"));
                    CU(cudaLibraryGetKernel(&v.kEmitProbe, v.lib, "pgpu_k_emit_probe"));
                }
            }
            if (v.gen.mode != KernelMode::COMPACT && p->plan->hasAgg && inputIsIntermediate(p->plan->step))
                CU(cudaLibraryGetKernel(&v.kGather, v.lib, "pgpu_k_gather"));
            if (v.gen.streamOk) {
                CU(cudaLibraryGetKernel(&v.kRunsCount, v.lib, "pgpu_k_runs_count"));
                CU(cudaLibraryGetKernel(&v.kRunsWrite, v.lib, "pgpu_k_runs_write"));
                CU(cudaLibraryGetKernel(&v.kRunsEmit, v.lib, "pgpu_k_runs_emit"));
                CU(cudaLibraryGetKernel(&v.kRunsToTable, v.lib, "pgpu_k_runs_to_table"));
                int nb = 0;
                if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, (const void*)v.kRunsWrite, kThreads, 0) != cudaSuccess || nb < 1) {
                    cudaGetLastError();
                    nb = 2;
                }
                v.streamBlocksPerSM = nb;
            }
            if (v.gen.sweepOk) {
                CU(cudaLibraryGetKernel(&v.kPartCount, v.lib, "pgpu_k_part_count"));
                CU(cudaLibraryGetKernel(&v.kPartScatter, v.lib, "pgpu_k_part_scatter"));
                CU(cudaLibraryGetKernel(&v.kPageSweep, v.lib, "pgpu_k_page_sweep"));
                CU(cudaFuncSetAttribute((const void*)v.kPartScatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)partScatterSmem(kPartMaxK, v.gen.sweepStageWidth)));
                if (v.gen.smemBytes > 48 * 1024) CU(cudaFuncSetAttribute((const void*)v.kPageSweep, cudaFuncAttributeMaxDynamicSharedMemorySize, v.gen.smemBytes));
                int nb = 0;
                if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, (const void*)v.kPageSweep, kThreads, (size_t)v.gen.smemBytes) != cudaSuccess || nb < 1) {
                    cudaGetLastError();
                    nb = 2;
                }
                v.sweepBlocksPerSM = nb;
            }
            if (v.gen.smemBytes > 48 * 1024) CU(cudaFuncSetAttribute((const void*)v.kPage, cudaFuncAttributeMaxDynamicSharedMemorySize, v.gen.smemBytes));
            if (v.gen.staged) {
                // ring depth from the opt-in shared memory of the device (227 KB on sm_100), at most 4 stages
                int optin = 0;
                CU(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, p->device));
                int stages = (int)std::min<long>(4, ((long)optin - 1024 - v.gen.stagedFixedSmem) / std::max(v.gen.stageBytes, 1));
                if (const char* e = getenv("PGPU_TUNE_STAGES")) stages = std::min(stages, atoi(e));
                if (stages >= 2) {
                    CU(cudaLibraryGetKernel(&v.kPageStaged, v.lib, "pgpu_k_page_staged"));
                    v.stagedStages = stages;
                    v.stagedSmem = (size_t)v.gen.stagedFixedSmem + (size_t)stages * v.gen.stageBytes;
                    CU(cudaFuncSetAttribute((const void*)v.kPageStaged, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)v.stagedSmem));
                }
            }
            int nb = 0;
            cudaError_t oe = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, (const void*)v.kPage, kThreads, (size_t)v.gen.smemBytes);
            if (oe != cudaSuccess || nb < 1) {
                cudaGetLastError();
                nb = 2;
            }
            v.blocksPerSM = nb;
        }
    } catch (const PlanError& e) {
        return fail(p, (pgpu_status)e.status, e.what());
    }
    auto ins = p->variants.emplace(key, std::move(v));
    *out = &ins.first->second;
    return PGPU_OK;
}

Scratch scratchOf(pgpu_pipeline* p) {
    Scratch s;
    s.err = p->dErr;
    s.partials = p->dPartials;
    s.ticket = p->dTicket;
    s.state = p->dState;
    s.tile_status = (uint64_t*)p->tileStatus.ptr;
    s.out_count = p->dOutCount;
    return s;
}

TableArgs tableOf(pgpu_pipeline* p) {
    TableArgs t;
    t.slots = p->dSlots;
    t.capacity = p->capacity;
    t.counters = p->dCounters;
    t.limit = ~0ull;  // no load-factor guard unless the caller sets one
    t.faillog = (uint64_t*)p->failLog.ptr;
    t.faillog_cap = 0;
    t.row_bias = 0;
    t.bucket_off = p->bucketOff;
    return t;
}

pgpu_status launch(pgpu_pipeline* p, cudaKernel_t k, int grid, int block, size_t smem, void** args) {
    if (grid < 1) grid = 1;
    if (p->timing) CU(cudaEventRecord(p->evA, p->stream));
    CU(cudaLaunchKernel((const void*)k, dim3((unsigned)grid), dim3((unsigned)block), args, smem, p->stream));
    if (p->timing) {
        CU(cudaEventRecord(p->evB, p->stream));
        CU(cudaEventSynchronize(p->evB));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, p->evA, p->evB));
        p->stats.device_ms += ms;
    }
    p->stats.kernel_launches++;
    return PGPU_OK;
}

pgpu_status checkDeviceErrors(pgpu_pipeline* p) {
    // caller has synchronised the stream
    uint32_t e = 0;
    CU(cudaMemcpyAsync(&p->hostScalars[3], p->dErr, 4, cudaMemcpyDeviceToHost, p->stream));
    CU(cudaStreamSynchronize(p->stream));
    e = (uint32_t)p->hostScalars[3];
    if (!e) return PGPU_OK;
    if (e & 2u) return fail(p, PGPU_DIVISION_BY_ZERO, "Division by zero");
    if (e & 1u) return fail(p, PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "numeric value out of range (integer or decimal overflow)");
    if (e & 32u) return fail(p, PGPU_INVALID_CAST_ARGUMENT, "Unable to cast double to bigint");
    if (e & 8u) return fail(p, PGPU_INVALID_ARGUMENT, "varchar(n) value longer than n bytes");
    if (e & 64u) return fail(p, PGPU_INVALID_ARGUMENT, "corrupt Parquet page: dictionary id out of range");
    if (e & 128u) return fail(p, PGPU_INVALID_STATE, "exchange: a peer's partial page did not arrive");
    if (e & 4u) return fail(p, PGPU_INSUFFICIENT_RESOURCES, "internal: hash table full");
    return fail(p, PGPU_INSUFFICIENT_RESOURCES, "output buffer overflow (gather exchange: more groups than an inbox slot holds)");
}

// ------------------------------------------------------------------------------------------- table management
pgpu_status allocTable(pgpu_pipeline* p, uint64_t cap, uint64_t** slotsOut) {
    size_t bytes = (size_t)cap * p->layout.SLOTW * 8;
    void* ptr = nullptr;
    cudaError_t e = cudaMalloc(&ptr, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(p, PGPU_INSUFFICIENT_RESOURCES, "cannot allocate a hash table of " + std::to_string(cap) + " slots (" + std::to_string(bytes >> 20) + " MiB)");
    }
    CU(cudaMemsetAsync(ptr, 0, bytes, p->stream));
    *slotsOut = (uint64_t*)ptr;
    return PGPU_OK;
}

// Look at the table's counters (synchronises): slots claimed, rows deferred. `exactGroups`: also learn the number of groups,
// which a bucketed table only knows by counting presence bits (one pass over the bucket headers).
pgpu_status refreshGroupCount(pgpu_pipeline* p, Variant* v = nullptr, bool exactGroups = false) {
    const bool buckets = p->tableMode == pgpu_pipeline::TABLE_BUCKETS;
    if (buckets && exactGroups && p->dSlots) {
        if (!v || !v->kCount) return fail(p, PGPU_CUDA_ERROR, "internal: group count without a bucket variant");
        CU(cudaMemsetAsync(p->dOutCount, 0, 8, p->stream));
        TableArgs t = tableOf(p);
        Scratch sc = scratchOf(p);
        void* args[] = {&t, &sc};
        int grid = (int)std::min<uint64_t>((p->capacity + kThreads - 1) / kThreads, (uint64_t)p->numSMs * 8);
        pgpu_status st = launch(p, v->kCount, grid, kThreads, 0, args);
        if (st) return st;
        CU(cudaMemcpyAsync(&p->hostScalars[0], p->dOutCount, 8, cudaMemcpyDeviceToHost, p->stream));
    }
    CU(cudaMemcpyAsync(&p->hostScalars[1], p->dCounters, 16, cudaMemcpyDeviceToHost, p->stream));
    CU(cudaStreamSynchronize(p->stream));
    p->slotsKnown = (int64_t)p->hostScalars[1];
    p->stats.table_slots_claimed = p->slotsKnown;
    p->stats.bucket_offset = (int64_t)p->bucketOff;
    p->groupsKnown = !buckets ? p->slotsKnown : exactGroups ? (int64_t)p->hostScalars[0] : p->slotsKnown * 8;  // upper bound
    p->rowsSinceKnown = 0;
    return PGPU_OK;
}

pgpu_status growTable(pgpu_pipeline* p, Variant* v, uint64_t newCap) {
    uint64_t* fresh = nullptr;
    pgpu_status st = allocTable(p, newCap, &fresh);
    if (st) return st;
    if (p->dSlots) CU(cudaMemsetAsync(p->dCounters, 0, 16, p->stream));  // groups are re-counted, deferred rows cleared
    if (p->dSlots && p->slotsKnown > 0) {
        TableArgs from = tableOf(p);
        TableArgs to = from;
        to.slots = fresh;
        to.capacity = newCap;
        Scratch sc = scratchOf(p);
        void* args[] = {&from, &to, &sc};
        int grid = (int)std::min<uint64_t>((from.capacity + kThreads - 1) / kThreads, (uint64_t)p->numSMs * 8);
        st = launch(p, v->kRehash, grid, kThreads, 0, args);
        if (st) return st;
        p->stats.table_rehashes++;
    }
    if (p->dSlots) {
        CU(cudaStreamSynchronize(p->stream));
        CU(cudaFree(p->dSlots));
    }
    p->dSlots = fresh;
    p->capacity = newCap;
    p->stats.table_capacity = (int64_t)newCap;
    return PGPU_OK;
}

// slots kept free below the load-factor limit for the end-of-kernel directory flush of every resident block
uint64_t flushReserve(pgpu_pipeline* p, const Variant* v) {
    if (!v->gen.useDir) return 0;
    return (uint64_t)p->numSMs * (uint64_t)v->blocksPerSM * (uint64_t)std::max(v->gen.dirCap * 3 / 4, 1);
}

// ------------------------------------------------------------------------------------------- page processing
struct DevPage {
    std::vector<ColArg> cols;
    int64_t rows = 0;
    int64_t slack = 0;  // rows of the same arrays that follow this slice (offsets past the end stay readable)
};

pgpu_status getVariant(pgpu_pipeline* p, const std::vector<ColSig>& sig, Variant** out);

// the variant of the same page signature whose kernels work on the bucketed table (callers hold variants that were looked up
// before the table format was decided: the rest of a chunked page, a pending super-page)
pgpu_status bucketVariantOf(pgpu_pipeline* p, Variant* v, Variant** out) {
    std::vector<ColSig> sig;
    for (auto& kv : p->variants)
        if (&kv.second == v) {
            for (size_t i = 0; i + 1 < kv.first.size(); i += 2) sig.push_back(ColSig{kv.first[i] - '0', kv.first[i + 1] == 'n'});
        }
    return getVariant(p, sig, out);
}

// First page of a high-cardinality plan that qualifies for the bucketed table: look at how its keys arrive. When most key
// changes stay inside a bucket of 8 neighbouring keys (l_orderkey: 7 of 8), 8 groups share one slot and one DRAM round trip;
// keys that jump around (shuffled keys, hash-partitioned partial results) keep the one-key-per-slot table.
// PGPU_TUNE_BUCKETS = 0 / 1 overrides the probe. *swapped = the bucket variant of the same page signature, if chosen.
pgpu_status decideTableMode(pgpu_pipeline* p, Variant* v, const DevPage& pg, Variant** swapped) {
    *swapped = nullptr;
    const GenResult& g = v->gen;
    bool buckets = false;
    const char* force = getenv("PGPU_TUNE_BUCKETS");
    const bool probe = force && !strcmp(force, "auto");
    if (!g.bucketsPossible || !force || !*force) {
        // default: one key per slot. Measured on GROUP BY l_orderkey at sf100 (profiles/r02_summary.md): the bucketed table cuts
        // the DRAM traffic per launch from 3.8x to 1.5x the input, but the kernel is bound by atomic round trips through L2, not
        // by bytes, and ends up slower (27.0 vs 24.4 ms per pass) — it stays behind PGPU_TUNE_BUCKETS = 1 (always) / auto (probe)
        buckets = false;
    } else if (!probe) {
        buckets = atoi(force) != 0;
    } else if (g.bucketKeyChannel >= 0 && pg.rows >= 1024) {
        const ColArg& ca = pg.cols[g.bucketKeyChannel];
        const Type& t = p->plan->input[g.bucketKeyChannel];
        if (ca.values && !ca.ids && !ca.dvalues && !t.isVarWidth() && (t.width() == 4 || t.width() == 8)) {
            unsigned long long* dOut = (unsigned long long*)p->dState;  // 64 words of scratch the hash path does not use
            CU(cudaMemsetAsync(dOut, 0, 80, p->stream));
            CU(keyClustering(p->stream, ca.values, t.width(), std::min<int64_t>(pg.rows, 1 << 16), dOut));
            CU(cudaMemcpyAsync(&p->hostScalars[4], dOut, 80, cudaMemcpyDeviceToHost, p->stream));
            CU(cudaStreamSynchronize(p->stream));
            CU(cudaMemsetAsync(dOut, 0, 80, p->stream));
            p->stats.kernel_launches++;
            const uint64_t changes = p->hostScalars[4];
            int best = 0;
            for (int d = 1; d < 8; d++)
                if (p->hostScalars[5 + d] > p->hostScalars[5 + best]) best = d;
            const uint64_t near = p->hostScalars[5 + best];
            buckets = changes > 0 && near * 2 >= changes;
            if (buckets) {
                p->bucketOff = (uint64_t)best;
                // a key change stays inside the bucket with probability near / changes: runs of 1 / (1 - that) keys per bucket
                p->bucketOccupancy = std::min(8.0, std::max(1.0, (double)changes / (double)std::max<uint64_t>(changes - near, 1)));
            }
        }
    }
    if (buckets && !probe) {
        if (const char* off = getenv("PGPU_TUNE_BUCKET_OFF")) p->bucketOff = (uint64_t)(atoi(off) & 7);
        p->bucketOccupancy = 4.0;
    }
    p->tableMode = buckets ? pgpu_pipeline::TABLE_BUCKETS : pgpu_pipeline::TABLE_PER_KEY;
    p->stats.table_bucketed = buckets ? 1 : 0;
    if (!buckets) return PGPU_OK;
    // same page signature, bucketed kernels; the slot layout changes with it (no table exists yet)
    p->haveLayout = false;
    return bucketVariantOf(p, v, swapped);
}

// Slots of a table that does not exist yet: BigintGroupByHash sizing (expectedSize / fill ratio, operator/BigintGroupByHash.java:78-101)
uint64_t initialTableCapacity(pgpu_pipeline* p, Variant* v) {
    const GenResult& g = v->gen;
    const uint64_t reserve = flushReserve(p, v);
    uint64_t cap = (uint64_t)((double)(p->tableSizeHint > 0 ? p->tableSizeHint : p->plan->expectedGroups) / 0.6) + 1024;
    if (g.buckets) cap = (uint64_t)((double)p->plan->expectedGroups / p->bucketOccupancy / 0.6) + 1024;  // keys per slot as the probe saw them
    cap = std::max<uint64_t>(cap, reserve * 4 + 4096);
    // key types that bound the number of groups (<= 1 M): size the table for the whole domain, once
    if (g.keyDomain > 0 && g.keyDomain <= (double)(1 << 20)) cap = std::max<uint64_t>(cap, (uint64_t)(g.keyDomain / 0.7) + reserve * 2 + 4096);
    return cap;
}

// the variant that buffered the runs (its merge / emit kernels share the run buffer's row format)
Variant* streamVariant(pgpu_pipeline* p) {
    for (auto& kv : p->variants)
        if (kv.second.kRunsEmit) return &kv.second;
    return nullptr;
}

// Streaming runs: the buffered rows go into the hash table (the key order broke, or a page arrives that the streaming kernel does
// not take); from here on the pipeline aggregates through the table.
pgpu_status streamToTable(pgpu_pipeline* p, Variant* v) {
    p->streamMode = pgpu_pipeline::STREAM_OFF;
    if (p->runRows == 0) return PGPU_OK;
    if (!v || !v->kRunsToTable) v = streamVariant(p);
    if (!v || !v->kRunsToTable) return fail(p, PGPU_CUDA_ERROR, "internal: buffered runs without a variant that can merge them");
    const uint64_t need = (uint64_t)((double)(p->slotsKnown + p->rowsSinceKnown + p->runRows) / 0.6) + 4096;
    pgpu_status st;
    if (!p->dSlots) {
        if ((st = growTable(p, v, std::max<uint64_t>(need, initialTableCapacity(p, v))))) return st;
    } else if (p->capacity < need) {
        if ((st = growTable(p, v, need))) return st;
    }
    TableArgs t = tableOf(p);
    Scratch sc = scratchOf(p);
    uint64_t* buf = (uint64_t*)p->runBuf.ptr;
    uint64_t n = (uint64_t)p->runRows;
    void* args[] = {&buf, &n, &t, &sc};
    int grid = (int)std::min<uint64_t>((n + kThreads - 1) / kThreads, (uint64_t)p->numSMs * 8);
    if ((st = launch(p, v->kRunsToTable, grid, kThreads, 0, args))) return st;
    p->rowsSinceKnown += p->runRows;
    p->runRows = 0;
    return PGPU_OK;
}

// sw: the page is a chunk of a partition-ordered batch (runSweep): the sweep kernel takes it
pgpu_status launchPageKernel(pgpu_pipeline* p, Variant* v, const DevPage& pg, const SweepArgs* sw = nullptr) {
    const GenResult& g = v->gen;
    if (g.mode == KernelMode::AGG_HASH && p->streamMode == pgpu_pipeline::STREAM_ON) {  // a page the streaming kernel does not take
        pgpu_status st = streamToTable(p, v);
        if (st) return st;
    }
    // kernel parameter block: PageArgs<NC> = ColArg[NC] + rows
    std::vector<unsigned char> pa(sizeof(ColArg) * g.NC + 8);
    memcpy(pa.data(), pg.cols.data(), sizeof(ColArg) * g.NC);
    memcpy(pa.data() + sizeof(ColArg) * g.NC, &pg.rows, 8);
    int64_t ntiles = (pg.rows + kTileRows - 1) / kTileRows;
    int grid = (int)std::min<int64_t>(ntiles, (int64_t)p->numSMs * v->blocksPerSM);
    Scratch sc = scratchOf(p);
    if (g.mode == KernelMode::AGG_GLOBAL) {
        size_t need = (size_t)grid * std::max(g.NW, 1) * 8;
        if (p->partialsCap < need) {
            if (p->dPartials) {
                CU(cudaStreamSynchronize(p->stream));
                CU(cudaFree(p->dPartials));
            }
            size_t cap = (size_t)p->numSMs * 32 * std::max(g.NW, 1) * 8;
            cap = std::max(cap, need);
            CU(cudaMalloc((void**)&p->dPartials, cap));
            p->partialsCap = cap;
            sc = scratchOf(p);
        }
        void* args[] = {pa.data(), &sc};
        return launch(p, v->kPage, grid, kThreads, 0, args);
    }
    if (g.mode == KernelMode::AGG_HASH && p->tableMode == pgpu_pipeline::TABLE_BUCKETS && !g.buckets) {
        Variant* swapped = nullptr;
        pgpu_status st = bucketVariantOf(p, v, &swapped);
        if (st) return st;
        if (!swapped->gen.buckets) return fail(p, PGPU_CUDA_ERROR, "internal: no bucketed variant for this page signature");
        return launchPageKernel(p, swapped, pg);
    }
    if (g.mode == KernelMode::AGG_HASH && !p->dSlots && p->tableMode == pgpu_pipeline::TABLE_UNDECIDED) {
        Variant* swapped = nullptr;
        pgpu_status st = decideTableMode(p, v, pg, &swapped);
        if (st) return st;
        if (swapped) return launchPageKernel(p, swapped, pg);
    }
    if (g.mode == KernelMode::AGG_HASH) {
        // Table sizing follows BigintGroupByHash (expectedSize / fill ratio, operator/BigintGroupByHash.java:78-101) and
        // grows on demand (:262-313). A chunk whose rows could push the table past its load-factor limit runs with a
        // deferred-row log: rows that would need a new slot beyond the limit are logged instead of inserted, the host then
        // grows the table (rehash) and replays exactly those rows. Nothing is dropped, nothing is over-allocated.
        const uint64_t reserve = flushReserve(p, v);
        if (!p->dSlots) {
            pgpu_status st = growTable(p, v, initialTableCapacity(p, v));
            if (st) return st;
        }
        const int64_t n = pg.rows;
        const uint64_t limit = (uint64_t)(0.75 * (double)p->capacity) - reserve;
        const bool bounded = g.keyDomain > 0 && (uint64_t)g.keyDomain <= limit;  // the table can hold every possible key
        const bool risky = !bounded && (uint64_t)(p->slotsKnown + p->rowsSinceKnown + n) > limit;
        if (risky) {
            pgpu_status st = ensureDev(p, p->failLog, (size_t)n * 8);
            if (st) return st;
        }
        TableArgs t = tableOf(p);
        t.limit = risky ? limit : ~0ull;  // the load-factor guard is only armed for chunks that could reach the limit
        t.faillog_cap = risky ? (uint64_t)n : 0;
        pgpu_status st;
        // staged variant (bulk-async ring in shared memory) for pages long enough to fill the ring on every SM and whose
        // arrays are 16-byte aligned (cp.async.bulk); short or unaligned pages take the direct-load kernel
        bool staged = !sw && v->kPageStaged && n >= (int64_t)g.stagedTile * p->numSMs;
        if (staged) {
            for (int c : g.usedChannels) {
                const ColArg& ca = pg.cols[c];
                if (((uintptr_t)ca.values | (uintptr_t)ca.nulls | (uintptr_t)ca.offsets | (uintptr_t)ca.ids) & 15) staged = false;
            }
            // the staged kernel decodes small dictionaries into shared memory once per block; larger ones take the direct kernel
            for (int c : g.dcacheChannels)
                if (pg.cols[c].dcount > g.dcacheCap) staged = false;
        }
        if (staged) {
            // whole tiles through the staged kernel, the ragged tail (and the 3 offsets a var-width tile reads past its end)
            // through the direct kernel: same table, same stream
            const int64_t nst = (n - (pg.slack >= 3 ? 0 : 3)) / g.stagedTile * g.stagedTile;
            std::vector<unsigned char> ps = pa;
            memcpy(ps.data() + sizeof(ColArg) * g.NC, &nst, 8);
            int nstages = v->stagedStages;
            void* args[] = {ps.data(), &t, &sc, &nstages};
            st = launch(p, v->kPageStaged, p->numSMs, g.stagedThreads, v->stagedSmem, args);
            if (st) return st;
            const int64_t nt = n - nst;
            if (nt > 0) {
                std::vector<ColArg> tc = pg.cols;
                for (int c : g.usedChannels) {
                    ColArg& ca = tc[c];
                    const Type& ty = p->plan->input[c];
                    if (ca.ids) ca.ids += nst;
                    if (ca.offsets) ca.offsets += nst;                                   // var-width: values stay (offsets are absolute)
                    else if (ca.values) ca.values = (const char*)ca.values + (size_t)nst * ty.width();
                    if (ca.nulls) ca.nulls += nst;
                }
                memcpy(ps.data(), tc.data(), sizeof(ColArg) * g.NC);
                memcpy(ps.data() + sizeof(ColArg) * g.NC, &nt, 8);
                TableArgs tt = t;
                tt.row_bias = (uint64_t)nst;  // deferred rows are logged with chunk-relative numbers (the replay runs on the chunk)
                int tgrid = (int)std::min<int64_t>((nt + kTileRows - 1) / kTileRows, (int64_t)p->numSMs * v->blocksPerSM);
                void* targs[] = {ps.data(), &tt, &sc};
                st = launch(p, v->kPage, tgrid, kThreads, (size_t)g.smemBytes, targs);
            }
        } else if (sw) {
            SweepArgs sa = *sw;
            CU(cudaMemsetAsync(sa.ticket, 0, 4, p->stream));
            void* args[] = {pa.data(), &t, &sc, &sa};
            cudaKernel_t k = v->kPageSweep;
            static const int64_t wt = [] { const char* e = getenv("PGPU_TUNE_SWEEP_WT"); int64_t v = e ? atoll(e) : 1024; return std::min<int64_t>(std::max<int64_t>(v / 128 * 128, 128), 4096); }();
            const int64_t nwarpTickets = (pg.rows + wt - 1) / wt;
            st = launch(p, k, (int)std::min<int64_t>((nwarpTickets + 7) / 8, (int64_t)p->numSMs * v->sweepBlocksPerSM), kThreads, (size_t)g.smemBytes, args);
        } else {
            void* args[] = {pa.data(), &t, &sc};
            st = launch(p, v->kPage, grid, kThreads, (size_t)g.smemBytes, args);
        }
        if (st) return st;
        p->rowsSinceKnown += n;
        if (!risky) return PGPU_OK;
        st = refreshGroupCount(p, v);  // synchronises; hostScalars[2] = deferred rows
        if (st) return st;
        uint64_t deferred = p->hostScalars[2];
        if (deferred == 0) return PGPU_OK;
        if (deferred > (uint64_t)n) return fail(p, PGPU_CUDA_ERROR, "internal: deferred-row log overflow");
        // every deferred row of a one-key-per-slot table may be a new group; deferred rows of a bucketed table share slots
        const double perSlot = g.buckets ? std::max(1.0, p->bucketOccupancy) * 2.0 : 1.0;
        uint64_t cap = std::max<uint64_t>(p->capacity * 2, (uint64_t)(((double)p->slotsKnown + (double)deferred / perSlot) / 0.6) + reserve * 4 + 4096);
        st = growTable(p, v, cap);  // rehash re-counts the groups and clears the deferred counter
        if (st) return st;
        TableArgs t2 = tableOf(p);  // limit = none: the grown table holds every deferred row even if all are new groups
        uint64_t* log = (uint64_t*)p->failLog.ptr;
        void* rargs[] = {pa.data(), &t2, &sc, &log, &deferred};
        int rgrid = (int)std::min<uint64_t>((deferred + kThreads - 1) / kThreads, (uint64_t)p->numSMs * 8);
        st = launch(p, v->kReplay, rgrid, kThreads, 0, rargs);
        if (st) return st;
        p->rowsSinceKnown += (int64_t)deferred;
        return PGPU_OK;
    }
    return fail(p, PGPU_INVALID_STATE, "internal: launchPageKernel in COMPACT mode");
}

// Streaming runs over a device-resident page (device header, "streaming runs"). *done = false when the page is left to the
// other paths: keys not in order (clustering probe of the first 65 536 keys, or seen to break earlier), small tables, short pages.
constexpr int64_t kStreamMinRows = 1ll << 20;
pgpu_status runStream(pgpu_pipeline* p, Variant* v, const DevPage& pg, bool* done) {
    *done = false;
    const GenResult& g = v->gen;
    if (!v->kRunsWrite || p->streamMode == pgpu_pipeline::STREAM_OFF || p->dSlots) return PGPU_OK;
    auto knob = [](const char* name, int64_t dflt) { const char* e = getenv(name); return e ? (int64_t)atoll(e) : dflt; };
    double runsPerRow = 0.5;
    if (p->streamMode == pgpu_pipeline::STREAM_UNDECIDED) {
        p->streamMode = pgpu_pipeline::STREAM_OFF;
        if (pg.rows < knob("PGPU_TUNE_STREAM_MIN_ROWS", kStreamMinRows)) return PGPU_OK;
        if ((size_t)initialTableCapacity(p, v) * g.SLOTW * 8 < (size_t)knob("PGPU_TUNE_STREAM_MIN_TABLE_MB", (int64_t)(kSweepMinTableBytes >> 20)) << 20) return PGPU_OK;
        if (p->tableMode == pgpu_pipeline::TABLE_BUCKETS || g.streamKeyChannel < 0) return PGPU_OK;
        if (const char* force = getenv("PGPU_TUNE_BUCKETS"))
            if (*force && strcmp(force, "0")) return PGPU_OK;  // the bucketed-table experiment was asked for: its probe decides, on the table path
        const ColArg& ca = pg.cols[g.streamKeyChannel];  // the first grouping field: in order it must be, whatever the others do
        const Type& t = p->plan->input[g.streamKeyChannel];
        if (!(ca.values && !ca.ids && !ca.dvalues && !t.isVarWidth() && (t.width() == 4 || t.width() == 8))) return PGPU_OK;
        unsigned long long* dOut = (unsigned long long*)p->dState;  // 64 words of scratch the hash path does not use
        const int64_t n = std::min<int64_t>(pg.rows, 1 << 16);
        CU(cudaMemsetAsync(dOut, 0, 80, p->stream));
        CU(keyClustering(p->stream, ca.values, t.width(), n, dOut));
        CU(cudaMemcpyAsync(&p->hostScalars[4], dOut, 80, cudaMemcpyDeviceToHost, p->stream));
        CU(cudaStreamSynchronize(p->stream));
        CU(cudaMemsetAsync(dOut, 0, 80, p->stream));
        p->stats.kernel_launches++;
        const uint64_t changes = p->hostScalars[4], descents = p->hostScalars[13];
        // in order, and in runs or at least dense enough that the buffer stays well below a hash table's size
        // (long groups belong to the table: its run compaction folds them per warp, and the emit kernel here walks a group serially)
        if ((descents != 0 || (changes + 1) * 32 < (uint64_t)n) && !knob("PGPU_TUNE_STREAM_FORCE", 0)) return PGPU_OK;
        runsPerRow = std::min(1.0, (double)(changes + 1) / (double)n * 1.3 + 0.02);
        p->streamMode = pgpu_pipeline::STREAM_ON;
    }
    if (!p->dViolations) CU(cudaMalloc((void**)&p->dViolations, 256));
    pgpu_status st;
    Scratch sc = scratchOf(p);
    const size_t rowBytes = (size_t)g.SLOTW * 8;
    // the buffer holds at least `needRows` rows (the rows it has are kept)
    auto ensureRunRows = [&](int64_t needRows, int64_t wantRows) -> pgpu_status {
        if (p->runBuf.cap >= (size_t)needRows * rowBytes) return PGPU_OK;
        DeviceBuf old = p->runBuf;
        p->runBuf = DeviceBuf();
        // (pages keep coming: grow by half at least, so that the rows kept are not copied once per page)
        pgpu_status e = ensureDev(p, p->runBuf, std::max((size_t)std::max(needRows, wantRows) * rowBytes, old.cap + old.cap / 2));
        if (e) return e;
        if (old.ptr) {
            if (p->runRows) CU(cudaMemcpyAsync(p->runBuf.ptr, old.ptr, (size_t)p->runRows * rowBytes, cudaMemcpyDeviceToDevice, p->stream));
            CU(cudaStreamSynchronize(p->stream));
            CU(cudaFree(old.ptr));
        }
        return PGPU_OK;
    };
    const int64_t chunkRows = (int64_t)1 << 30;  // (the unit counts of a launch stay below 2^32)
    for (int64_t r0 = 0; r0 < pg.rows; r0 += chunkRows) {
        const int64_t n = std::min(chunkRows, pg.rows - r0);
        std::vector<ColArg> cols = pg.cols;
        for (int c : g.usedChannels) {
            const Type& t = p->plan->input[c];
            ColArg& ca = cols[c];
            if (ca.ids) ca.ids += r0;
            if (ca.offsets) ca.offsets += r0;
            else if (ca.values) ca.values = (const char*)ca.values + (size_t)r0 * (t.isVarWidth() ? 16 : t.width());
            if (ca.nulls) ca.nulls += r0;
        }
        std::vector<unsigned char> pa(sizeof(ColArg) * g.NC + 8);
        memcpy(pa.data(), cols.data(), sizeof(ColArg) * g.NC);
        memcpy(pa.data() + sizeof(ColArg) * g.NC, &n, 8);
        const int grid = p->numSMs * v->streamBlocksPerSM;
        {
            // count the runs per warp unit, scan, size the buffer exactly, write
            const int64_t unitRows = 32 * 4 * g.streamSub, units = (n + unitRows - 1) / unitRows;
            if ((st = ensureDev(p, p->runCounts, (size_t)units * 4))) return st;
            if ((st = ensureDev(p, p->runOffsets, (size_t)(units + 1024) * 8))) return st;
            uint32_t* counts = (uint32_t*)p->runCounts.ptr;
            unsigned long long* offsets = (unsigned long long*)p->runOffsets.ptr;
            {
                void* args[] = {pa.data(), &counts, &sc};
                if ((st = launch(p, v->kRunsCount, (int)std::min<int64_t>((units + 7) / 8, grid), kThreads, 0, args))) return st;
            }
            CU(runsScan(p->stream, counts, units, offsets, offsets + units, p->numSMs * 4, (unsigned long long*)p->dOutCount));
            p->stats.kernel_launches += 2;
            CU(cudaMemcpyAsync(&p->hostScalars[0], p->dOutCount, 8, cudaMemcpyDeviceToHost, p->stream));
            CU(cudaStreamSynchronize(p->stream));
            const int64_t live = (int64_t)p->hostScalars[0];
            if ((st = ensureRunRows(p->runRows + live, p->runRows + live + (int64_t)((double)(pg.rows - r0 - n) * runsPerRow)))) return st;
            RunArgs ra{(uint64_t*)p->runBuf.ptr, (uint64_t)(p->runBuf.cap / rowBytes), (uint64_t)p->runRows};
            const unsigned long long* coffsets = offsets;
            void* args[] = {pa.data(), &coffsets, &ra, &sc};
            if ((st = launch(p, v->kRunsWrite, (int)std::min<int64_t>((units + 7) / 8, grid), kThreads, 0, args))) return st;
            p->runRows += live;
        }
    }
    p->stats.streamed_runs = p->runRows;
    *done = true;
    return PGPU_OK;
}

// Partition sweep (device header, "partition sweep") over a long device-resident page: count rows per table slice, scatter the
// used channels' arrays into slice order, run the page kernel's sweep variant over the reordered rows. *swept = false when the
// page is left to the ordinary path (table not much larger than L2, no memory for the reordered copy, bucketed table).
pgpu_status runSweep(pgpu_pipeline* p, Variant* v, const DevPage& pg, bool* swept) {
    *swept = false;
    const GenResult& g = v->gen;
    const Plan& plan = *p->plan;
    if (!v->kPageSweep) return PGPU_OK;
    if (p->tableMode == pgpu_pipeline::TABLE_BUCKETS) return PGPU_OK;
    if (p->tableMode == pgpu_pipeline::TABLE_UNDECIDED) {
        if (p->dSlots) return PGPU_OK;
        DevPage first = pg;
        first.rows = std::min<int64_t>(pg.rows, kHashChunkRows);
        Variant* swapped = nullptr;
        pgpu_status st = decideTableMode(p, v, first, &swapped);
        if (st) return st;
        if (swapped || p->tableMode != pgpu_pipeline::TABLE_PER_KEY) return PGPU_OK;
    }
    // test / experiment knobs: PGPU_TUNE_SWEEP_MIN_ROWS, PGPU_TUNE_SWEEP_MIN_TABLE_MB, PGPU_TUNE_SWEEP_BATCH_ROWS, PGPU_TUNE_SWEEP_K
    auto knob = [](const char* name, int64_t dflt) { const char* e = getenv(name); return e ? (int64_t)atoll(e) : dflt; };
    const int64_t minRows = knob("PGPU_TUNE_SWEEP_MIN_ROWS", kSweepMinRows);
    if (pg.rows < minRows) return PGPU_OK;
    const uint64_t cap = p->dSlots ? p->capacity : initialTableCapacity(p, v);
    const size_t tableBytes = (size_t)cap * g.SLOTW * 8;
    if (tableBytes < (size_t)knob("PGPU_TUNE_SWEEP_MIN_TABLE_MB", (int64_t)(kSweepMinTableBytes >> 20)) << 20) return PGPU_OK;
    // Keys that arrive in runs or in key order are cheap as they are (run compaction makes one table update per run, neighbouring
    // keys share table lines): measured on sf100 GROUP BY l_orderkey, runs of 4 rows, 19.7 ms on the ordinary path against 25 ms
    // here; keys in no order 57 ms against 38 ms. The clustering probe of the bucketed table looks at the first 65536 keys.
    if (!knob("PGPU_TUNE_SWEEP_FORCE", 0) && g.bucketKeyChannel >= 0 && pg.rows >= 1024) {
        const ColArg& ca = pg.cols[g.bucketKeyChannel];
        const Type& t = plan.input[g.bucketKeyChannel];
        if (ca.values && !ca.ids && !ca.dvalues && !t.isVarWidth() && (t.width() == 4 || t.width() == 8)) {
            unsigned long long* dOut = (unsigned long long*)p->dState;  // 64 words of scratch the hash path does not use
            const int64_t n = std::min<int64_t>(pg.rows, 1 << 16);
            CU(cudaMemsetAsync(dOut, 0, 80, p->stream));
            CU(keyClustering(p->stream, ca.values, t.width(), n, dOut));
            CU(cudaMemcpyAsync(&p->hostScalars[4], dOut, 80, cudaMemcpyDeviceToHost, p->stream));
            CU(cudaStreamSynchronize(p->stream));
            CU(cudaMemsetAsync(dOut, 0, 80, p->stream));
            p->stats.kernel_launches++;
            uint64_t near = 0;
            for (int d = 0; d < 8; d++) near = std::max<uint64_t>(near, p->hostScalars[5 + d]);
            const double jumps = (double)p->hostScalars[4] - 0.75 * (double)near;  // key changes that land on another table line
            if (jumps < 0.45 * (double)n) return PGPU_OK;
        }
    }
    // the reordered copy: one array per moved array of the used channels, each 256-byte aligned
    struct Moved {
        int channel;
        int kind;   // 0 values, 1 nulls, 2 ids
        int width;
        size_t off;
    };
    std::vector<Moved> moved;
    size_t perRow = 0;
    for (int c : g.usedChannels) {
        const Type& t = plan.input[c];
        const ColArg& ca = pg.cols[c];
        if (ca.ids) {
            moved.push_back({c, 2, 4, 0});
            perRow += 4;
        } else if (ca.values) {  // RLE blocks have neither
            const int w = t.isVarWidth() ? 16 : t.width();
            moved.push_back({c, 0, w, 0});
            perRow += w;
            if (ca.nulls) {
                moved.push_back({c, 1, 1, 0});
                perRow += 1;
            }
        }
    }
    if (perRow == 0) return PGPU_OK;
    static const bool trace = getenv("PGPU_TRACE_SWEEP") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = trace ? now() : 0;
    // (cudaMemGetInfo costs 0.2 - 2 ms, now and then 75 ms: only asked when the buffer of an earlier batch does not already fit the page)
    size_t freeB = 0, totalB = 0;
    const bool fits = p->dSlots && p->sweepBuf.cap >= (size_t)pg.rows * perRow + moved.size() * 256;
    if (!fits) CU(cudaMemGetInfo(&freeB, &totalB));
    const double t1 = trace ? now() : 0;
    const size_t budget = p->sweepBuf.cap + (size_t)((double)freeB * 0.6);
    const size_t table0 = p->dSlots ? 0 : tableBytes;  // the table is still to be allocated
    if (budget < table0 + (size_t)minRows * perRow) return PGPU_OK;
    int64_t batchRows = std::min<int64_t>(pg.rows, (int64_t)((budget - table0) / perRow));
    batchRows = std::min<int64_t>(batchRows, knob("PGPU_TUNE_SWEEP_BATCH_ROWS", batchRows));
    if (batchRows < std::min<int64_t>(minRows, pg.rows)) return PGPU_OK;
    int K = (int)std::min<size_t>(kPartMaxK, std::max<size_t>(8, (tableBytes + kSweepSliceBytes - 1) / kSweepSliceBytes));
    K = (int)std::min<int64_t>(kPartMaxK, std::max<int64_t>(1, knob("PGPU_TUNE_SWEEP_K", K)));
    pgpu_status st;
    {
        size_t need = 0;
        for (auto& m : moved) {
            m.off = need;
            need += align256((size_t)batchRows * m.width);
        }
        if ((st = ensureDev(p, p->sweepBuf, need))) return st;
        if ((st = ensureDev(p, p->sweepMeta, (size_t)kPartMaxK * 4 + 256 + (size_t)kPartMaxK * 8 + (size_t)(kPartMaxK + 1) * 8))) return st;
    }
    uint32_t* hist = (uint32_t*)p->sweepMeta.ptr;
    uint32_t* ticket = hist + kPartMaxK;
    uint64_t* cursor = (uint64_t*)((char*)p->sweepMeta.ptr + (size_t)kPartMaxK * 4 + 256);
    uint64_t* start = cursor + kPartMaxK;
    Scratch sc = scratchOf(p);
    for (int64_t b0 = 0; b0 < pg.rows; b0 += batchRows) {
        const int64_t bn = std::min(batchRows, pg.rows - b0);
        // input slice [b0, b0 + bn) and the reordered page over sweepBuf
        std::vector<ColArg> in = pg.cols, out = pg.cols;
        for (int c : g.usedChannels) {
            const Type& t = plan.input[c];
            ColArg& ca = in[c];
            if (ca.ids) ca.ids += b0;
            else if (ca.values) ca.values = (const char*)ca.values + (size_t)b0 * (t.isVarWidth() ? 16 : t.width());
            if (ca.nulls) ca.nulls += b0;
        }
        for (auto& m : moved) {
            char* base = (char*)p->sweepBuf.ptr + m.off;
            if (m.kind == 0) out[m.channel].values = base;
            else if (m.kind == 1) out[m.channel].nulls = (const uint8_t*)base;
            else out[m.channel].ids = (const int32_t*)base;
        }
        std::vector<unsigned char> pin(sizeof(ColArg) * g.NC + 8), pout(sizeof(ColArg) * g.NC + 8);
        memcpy(pin.data(), in.data(), sizeof(ColArg) * g.NC);
        memcpy(pin.data() + sizeof(ColArg) * g.NC, &bn, 8);
        memcpy(pout.data(), out.data(), sizeof(ColArg) * g.NC);
        memcpy(pout.data() + sizeof(ColArg) * g.NC, &bn, 8);
        PartArgs pa{hist, cursor, (uint32_t)K, 0};
        CU(cudaMemsetAsync(hist, 0, (size_t)kPartMaxK * 4, p->stream));
        {
            void* args[] = {pin.data(), &pa, &sc};
            const int64_t ntiles = (bn + kTileRows - 1) / kTileRows;
            if ((st = launch(p, v->kPartCount, (int)std::min<int64_t>(ntiles, (int64_t)p->numSMs * 8), kThreads, 0, args))) return st;
        }
        CU(partitionScan(p->stream, hist, (unsigned long long*)cursor, (unsigned long long*)start, K));
        {
            if (v->scatterBlocksPerSM == 1) {
                int nb = 0;
                if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, (const void*)v->kPartScatter, kPartThreads, partScatterSmem(K, g.sweepStageWidth)) != cudaSuccess || nb < 1) {
                    cudaGetLastError();
                    nb = 1;
                }
                v->scatterBlocksPerSM = nb;
            }
            void* args[] = {pin.data(), pout.data(), &pa};
            const int64_t ntiles = (bn + kPartTile - 1) / kPartTile;
            if ((st = launch(p, v->kPartScatter, (int)std::min<int64_t>(ntiles, (int64_t)p->numSMs * v->scatterBlocksPerSM), kPartThreads, partScatterSmem(K, g.sweepStageWidth), args))) return st;
        }
        const double t2 = trace ? now() : 0;
        CU(cudaMemcpyAsync(&p->hostScalars[13], start + K, 8, cudaMemcpyDeviceToHost, p->stream));
        CU(cudaStreamSynchronize(p->stream));
        const double t3 = trace ? now() : 0;
        const int64_t passed = (int64_t)p->hostScalars[13];  // rows that pass the filter
        if (passed < 0 || passed > bn) return fail(p, PGPU_CUDA_ERROR, "internal: partition sweep row count");
        p->stats.partition_sweeps++;
        SweepArgs sw{start, ticket, 0, (uint32_t)K, 0};
        for (int64_t r0 = 0; r0 < passed; r0 += kHashChunkRows) {
            DevPage chunk;
            chunk.rows = std::min(kHashChunkRows, passed - r0);
            chunk.slack = passed - (r0 + chunk.rows);
            chunk.cols = out;
            for (auto& m : moved) {
                ColArg& ca = chunk.cols[m.channel];
                if (m.kind == 0) ca.values = (const char*)ca.values + (size_t)r0 * m.width;
                else if (m.kind == 1) ca.nulls += r0;
                else ca.ids += r0;
            }
            sw.row0 = r0;
            if ((st = launchPageKernel(p, v, chunk, &sw))) return st;
        }
        if (trace)
            fprintf(stderr, "[pgpu sweep] rows %lld passed %lld K %d free %.1f GB: memGetInfo %.2f ms, alloc + count/scatter launches %.2f ms, wait %.2f ms, sweep launches %.2f ms\n",
                    (long long)bn, (long long)passed, K, freeB / 1e9, t1 - t0, t2 - t1, t3 - t2, now() - t3);
    }
    *swept = true;
    return PGPU_OK;
}

// keepRows > 0: the first keepRows rows of every output column survive a reallocation (buffered small pages)
pgpu_status ensureOutBuffers(pgpu_pipeline* p, int64_t rows, int64_t keepRows = 0) {
    const GenResult& g = p->layout;
    if ((int)p->outDev.size() != g.NO) {
        p->outDev.resize(g.NO);
        p->outDevNulls.resize(g.NO);
        p->outHost.resize(g.NO);
        p->outHostNulls.resize(g.NO);
        p->outHostAux.resize(g.NO);
    }
    int64_t r = std::max<int64_t>(rows, 1);
    if (p->outArena.ptr && p->outCapacityRows >= r && (int)p->outOffV.size() == g.NO) return PGPU_OK;  // layout stays
    DeviceBuf old;
    std::vector<size_t> oldOffV, oldOffN;
    if (keepRows > 0 && p->outArena.ptr) {
        r = std::max<int64_t>(r, 2 * p->outCapacityRows);
        old = p->outArena;
        p->outArena = DeviceBuf();
        oldOffV = p->outOffV;
        oldOffN = p->outOffN;
    }
    p->outOffV.assign(g.NO, 0);
    p->outOffN.assign(g.NO, 0);
    size_t need = 0;
    for (int c = 0; c < g.NO; c++) {
        p->outOffV[c] = need;
        need += align256((size_t)r * std::max(g.out[c].stride, 1));
        p->outOffN[c] = need;
        need += align256((size_t)r);
    }
    pgpu_status st = ensureDev(p, p->outArena, need);
    if (st) return st;
    p->outLayoutBytes = need;
    for (int c = 0; c < g.NO; c++) {
        p->outDev[c].ptr = (char*)p->outArena.ptr + p->outOffV[c];
        p->outDev[c].cap = (size_t)r * std::max(g.out[c].stride, 1);
        p->outDevNulls[c].ptr = (char*)p->outArena.ptr + p->outOffN[c];
        p->outDevNulls[c].cap = (size_t)r;
    }
    p->outCapacityRows = r;
    if (old.ptr) {
        for (int c = 0; c < g.NO; c++) {
            CU(cudaMemcpyAsync(p->outDev[c].ptr, (char*)old.ptr + oldOffV[c], (size_t)keepRows * std::max(g.out[c].stride, 1), cudaMemcpyDeviceToDevice, p->stream));
            CU(cudaMemcpyAsync(p->outDevNulls[c].ptr, (char*)old.ptr + oldOffN[c], (size_t)keepRows, cudaMemcpyDeviceToDevice, p->stream));
        }
        CU(cudaStreamSynchronize(p->stream));
        CU(cudaFree(old.ptr));
    }
    return PGPU_OK;
}

// baseRows: the kernel's row 0 lands at this row of the output arena
std::vector<unsigned char> outArgsBlock(pgpu_pipeline* p, int64_t baseRows = 0) {
    const GenResult& g = p->layout;
    std::vector<unsigned char> oa(sizeof(OutCol) * g.NO + 8);
    for (int c = 0; c < g.NO; c++) {
        OutCol oc{(char*)p->outDev[c].ptr + (size_t)baseRows * std::max(g.out[c].stride, 1), (uint8_t*)p->outDevNulls[c].ptr + baseRows};
        memcpy(oa.data() + sizeof(OutCol) * c, &oc, sizeof oc);
    }
    int64_t cap = p->outCapacityRows - baseRows;
    memcpy(oa.data() + sizeof(OutCol) * g.NO, &cap, 8);
    return oa;
}

pgpu_status runCompact(pgpu_pipeline* p, Variant* v, const DevPage& pg) {
    const GenResult& g = v->gen;
    const int64_t base = p->mergeRows;  // buffered small pages stay in front (MergePages)
    pgpu_status st = ensureOutBuffers(p, base + pg.rows, base);
    if (st) return st;
    int64_t ntiles = (pg.rows + kTileRows - 1) / kTileRows;
    st = ensureDev(p, p->tileStatus, (size_t)ntiles * 8);
    if (st) return st;
    CU(cudaMemsetAsync(p->tileStatus.ptr, 0, (size_t)ntiles * 8, p->stream));
    CU(cudaMemsetAsync(p->dTicket, 0, 4, p->stream));
    CU(cudaMemsetAsync(p->dOutCount, 0, 8, p->stream));
    std::vector<unsigned char> pa(sizeof(ColArg) * g.NC + 8);
    memcpy(pa.data(), pg.cols.data(), sizeof(ColArg) * g.NC);
    memcpy(pa.data() + sizeof(ColArg) * g.NC, &pg.rows, 8);
    std::vector<unsigned char> oa = outArgsBlock(p, base);
    Scratch sc = scratchOf(p);
    int grid = (int)std::min<int64_t>(ntiles, (int64_t)p->numSMs * v->blocksPerSM);
    void* args[] = {pa.data(), oa.data(), &sc};
    st = launch(p, v->kPage, grid, kThreads, 0, args);
    if (st) return st;
    CU(cudaMemcpyAsync(&p->hostScalars[0], p->dOutCount, 8, cudaMemcpyDeviceToHost, p->stream));
    p->outputReady = true;
    p->compactPending = true;  // the row count is read after the sync in getOutput
    return PGPU_OK;
}

// slices a device-resident page into row ranges and runs the aggregation kernels over them
pgpu_status processDevicePage(pgpu_pipeline* p, Variant* v, const pgpu_column* cols, int64_t rows) {
    const GenResult& g = v->gen;
    const Plan& plan = *p->plan;
    auto colArgsFor = [&](int64_t r0, std::vector<ColArg>& out) {
        out.assign(g.NC, ColArg{});
        for (int c : g.usedChannels) {
            const pgpu_column& col = cols[c];
            const Type& t = plan.input[c];
            ColArg a{};
            if (col.encoding == PGPU_ENC_VARCELL) {
                a.values = (const char*)col.values + (size_t)r0 * 16;
                a.nulls = col.nulls ? col.nulls + r0 : nullptr;
            } else if (col.encoding == PGPU_ENC_PLAIN) {
                if (t.isVarWidth()) {
                    a.values = col.values;
                    a.offsets = col.offsets + r0;
                } else {
                    a.values = (const char*)col.values + (size_t)r0 * t.width();
                }
                a.nulls = col.nulls ? col.nulls + r0 : nullptr;
            } else {
                const pgpu_column* d = col.dictionary;
                if (col.encoding == PGPU_ENC_DICTIONARY) a.ids = col.ids + r0;
                a.dvalues = d->values;
                a.dnulls = d->nulls;
                a.doffsets = d->offsets;
                a.dcount = (int32_t)std::min<int64_t>(d->position_count, INT32_MAX);
            }
            out[c] = a;
        }
    };
    if (g.mode == KernelMode::COMPACT) {
        DevPage pg;
        pg.rows = rows;
        colArgsFor(0, pg.cols);
        return runCompact(p, v, pg);
    }
    int64_t chunk = g.mode == KernelMode::AGG_HASH ? kHashChunkRows : rows;
    if (v->kPageStaged) chunk = chunk / g.stagedTile * g.stagedTile;  // inner chunks are whole tiles: only the page's end is ragged
    if (g.mode == KernelMode::AGG_HASH && v->kRunsWrite && p->streamMode != pgpu_pipeline::STREAM_OFF && !p->dSlots) {
        DevPage whole;
        whole.rows = rows;
        colArgsFor(0, whole.cols);
        bool done = false;
        pgpu_status st = runStream(p, v, whole, &done);
        if (st) return st;
        if (done) return PGPU_OK;
    }
    for (int64_t r0 = 0; r0 < rows; r0 += chunk) {
        if (v->kPageSweep) {
            // a long device-resident page and a table far larger than L2: reorder the rest of the page by table slice first
            DevPage rest;
            rest.rows = rows - r0;
            colArgsFor(r0, rest.cols);
            bool swept = false;
            pgpu_status st = runSweep(p, v, rest, &swept);
            if (st) return st;
            if (swept) return PGPU_OK;
        }
        int64_t n = std::min(chunk, rows - r0);
        DevPage pg;
        pg.rows = n;
        pg.slack = rows - (r0 + n);
        colArgsFor(r0, pg.cols);
        pgpu_status st = launchPageKernel(p, v, pg);
        if (st) return st;
    }
    return PGPU_OK;
}

pgpu_status validatePage(pgpu_pipeline* p, const pgpu_column* cols, int32_t ncols, int64_t rows, std::vector<ColSig>& sig) {
    const Plan& plan = *p->plan;
    if (!cols || ncols != (int)plan.input.size()) return fail(p, PGPU_INVALID_ARGUMENT, "page has " + std::to_string(ncols) + " channels, plan expects " + std::to_string(plan.input.size()));
    if (rows < 0) return fail(p, PGPU_INVALID_ARGUMENT, "negative position count");
    sig.resize(ncols);
    for (int c = 0; c < ncols; c++) {
        const pgpu_column& col = cols[c];
        const Type& t = plan.input[c];
        if (col.type_kind != (int)t.kind || (t.kind == TypeKind::DECIMAL && (col.precision != t.precision || col.scale != t.scale)))
            return fail(p, PGPU_INVALID_ARGUMENT, "channel " + std::to_string(c) + " type does not match the plan (" + t.str() + ")");
        if (col.position_count != rows) return fail(p, PGPU_INVALID_ARGUMENT, "channel " + std::to_string(c) + " position count differs from the page");
        sig[c].encoding = col.encoding;
        if (col.encoding == PGPU_ENC_PLAIN) {
            sig[c].nullable = col.nulls != nullptr;
            if (rows > 0 && !col.values && !(t.isVarWidth())) return fail(p, PGPU_INVALID_ARGUMENT, "channel " + std::to_string(c) + " has no values");
            if (t.isVarWidth() && !col.offsets) return fail(p, PGPU_INVALID_ARGUMENT, "var-width channel " + std::to_string(c) + " has no offsets");
        } else if (col.encoding == PGPU_ENC_VARCELL) {
            if (t.kind != TypeKind::VARCHAR || t.precision < 1 || t.precision > 15) return fail(p, PGPU_INVALID_ARGUMENT, "16-byte cell encoding is for VARCHAR(<=15) channels (channel " + std::to_string(c) + ")");
            if (rows > 0 && !col.values) return fail(p, PGPU_INVALID_ARGUMENT, "channel " + std::to_string(c) + " has no values");
            sig[c].nullable = col.nulls != nullptr;
        } else if (col.encoding == PGPU_ENC_DICTIONARY || col.encoding == PGPU_ENC_RLE) {
            const pgpu_column* d = col.dictionary;
            if (!d || d->encoding != PGPU_ENC_PLAIN) return fail(p, PGPU_NOT_SUPPORTED, "not supported on the GPU path: nested dictionary/RLE encodings (channel " + std::to_string(c) + ")");
            if (col.encoding == PGPU_ENC_DICTIONARY && !col.ids && rows > 0) return fail(p, PGPU_INVALID_ARGUMENT, "dictionary channel without ids");
            if (col.encoding == PGPU_ENC_RLE && d->position_count != 1) return fail(p, PGPU_INVALID_ARGUMENT, "RLE value block must have exactly one position");
            sig[c].nullable = d->nulls != nullptr;
        } else {
            return fail(p, PGPU_INVALID_ARGUMENT, "unknown block encoding");
        }
    }
    return PGPU_OK;
}


// Both halves of a staging slot hold at least `bytes` (the pinned half only when the library does the staging copy).
pgpu_status ensureSlot(pgpu_pipeline* p, StageSlot& slot, size_t bytes, bool needHost) {
    pgpu_status st = ensureDev(p, slot.dev, bytes);
    if (st) return st;
    if (needHost && slot.hostCap < bytes) {
        if (slot.host) CU(cudaFreeHost(slot.host));
        slot.host = nullptr;
        slot.hostCap = 0;
        // PGPU_STAGE_WC=1: write-combined pinned memory (the CPU only writes a slot, with non-temporal stores; the copy engine reads
        // it without snooping the caches). Experiment knob: the dictionary comparison of the appender reads the slot back.
        static const bool wc = getenv("PGPU_STAGE_WC") && atoi(getenv("PGPU_STAGE_WC")) != 0;
        if (wc) CU(cudaHostAlloc(&slot.host, bytes, cudaHostAllocWriteCombined));
        else CU(cudaMallocHost(&slot.host, bytes));
        slot.hostCap = bytes;
    }
    return PGPU_OK;
}

// The slot the next super-page / chunk goes into: waits until the kernel that read its previous contents has finished.
pgpu_status acquireSlot(pgpu_pipeline* p, StageSlot** out) {
    StageSlot& slot = p->slots[p->nextSlot];
    if (slot.inFlight) {
        CU(cudaEventSynchronize(slot.done));
        slot.inFlight = false;
    }
    *out = &slot;
    return PGPU_OK;
}

pgpu_status releaseSlot(pgpu_pipeline* p, StageSlot& slot) {
    if (!slot.done) CU(cudaEventCreateWithFlags(&slot.done, cudaEventDisableTiming));
    CU(cudaEventRecord(slot.done, p->stream));
    slot.inFlight = true;
    p->nextSlot = (p->nextSlot + 1) % 3;
    return PGPU_OK;
}

// Stage rows [r0, r0+n) of a host page into a ring slot and return device-side column descriptors (pages the coalescing
// appender below does not take: caller-pinned pages, filter + project pages, very large dictionaries).
pgpu_status stageChunk(pgpu_pipeline* p, Variant* v, const pgpu_column* cols, int64_t r0, int64_t n, bool pinned, StageSlot& slot, std::vector<ColArg>& out) {
    const GenResult& g = v->gen;
    const Plan& plan = *p->plan;
    struct Piece {
        const void* src;
        size_t bytes;
        size_t off;
    };
    std::vector<Piece> pieces;
    size_t total = 0;
    auto add = [&](const void* src, size_t bytes) -> size_t {
        size_t off = total;
        pieces.push_back({src, bytes, off});
        total += align256(std::max<size_t>(bytes, 1));
        return off;
    };
    struct ColOff {
        size_t values = 0, nulls = 0, offsets = 0, ids = 0, dvalues = 0, dnulls = 0, doffsets = 0;
        bool hasNulls = false, hasDnulls = false;
        int64_t byteBase = 0;
    };
    std::vector<ColOff> offs(g.NC);
    for (int c : g.usedChannels) {
        const pgpu_column& col = cols[c];
        const Type& t = plan.input[c];
        ColOff& o = offs[c];
        if (col.encoding == PGPU_ENC_PLAIN) {
            if (t.isVarWidth()) {
                // the slice starts at a 16-byte aligned byte offset so that device pointer + offset keeps the alignment the
                // bulk-copy path needs
                int32_t b0 = col.offsets[r0] & ~15, b1 = col.offsets[r0 + n];
                o.byteBase = b0;
                o.offsets = add(col.offsets + r0, (size_t)(n + 1) * 4);
                o.values = add((const char*)col.values + b0, (size_t)(b1 - b0));
            } else {
                o.values = add((const char*)col.values + (size_t)r0 * t.width(), (size_t)n * t.width());
            }
            if (col.nulls) {
                o.hasNulls = true;
                o.nulls = add(col.nulls + r0, (size_t)n);
            }
        } else {
            const pgpu_column* d = col.dictionary;
            if (col.encoding == PGPU_ENC_DICTIONARY) o.ids = add(col.ids + r0, (size_t)n * 4);
            int64_t dn = d->position_count;
            if (t.isVarWidth()) {
                o.doffsets = add(d->offsets, (size_t)(dn + 1) * 4);
                o.dvalues = add(d->values, (size_t)d->offsets[dn]);
            } else {
                o.dvalues = add(d->values, (size_t)dn * t.width());
            }
            if (d->nulls) {
                o.hasDnulls = true;
                o.dnulls = add(d->nulls, (size_t)dn);
            }
        }
    }
    pgpu_status st = ensureSlot(p, slot, total, !pinned);
    if (st) return st;
    char* dbase = (char*)slot.dev.ptr;
    if (pinned) {
        for (auto& pc : pieces)
            if (pc.bytes) CU(cudaMemcpyAsync(dbase + pc.off, pc.src, pc.bytes, cudaMemcpyHostToDevice, p->stream));
    } else {
        std::vector<CopyTask> tasks;
        for (auto& pc : pieces)
            if (pc.bytes) tasks.push_back(CopyTask{CopyTask::BYTES, (char*)slot.host + pc.off, pc.src, pc.bytes, 0});
        hostCopy(tasks);
        CU(cudaMemcpyAsync(dbase, slot.host, total, cudaMemcpyHostToDevice, p->stream));
    }
    for (auto& pc : pieces) p->stats.h2d_bytes += (int64_t)pc.bytes;
    out.assign(g.NC, ColArg{});
    for (int c : g.usedChannels) {
        const pgpu_column& col = cols[c];
        const Type& t = plan.input[c];
        const ColOff& o = offs[c];
        ColArg a{};
        if (col.encoding == PGPU_ENC_PLAIN) {
            if (t.isVarWidth()) {
                a.offsets = (const int32_t*)(dbase + o.offsets);
                a.values = dbase + o.values - o.byteBase;  // offsets stay absolute
            } else {
                a.values = dbase + o.values;
            }
            if (o.hasNulls) a.nulls = (const uint8_t*)(dbase + o.nulls);
        } else {
            if (col.encoding == PGPU_ENC_DICTIONARY) a.ids = (const int32_t*)(dbase + o.ids);
            a.dvalues = dbase + o.dvalues;
            a.dcount = (int32_t)std::min<int64_t>(col.dictionary->position_count, INT32_MAX);
            if (t.isVarWidth()) a.doffsets = (const int32_t*)(dbase + o.doffsets);
            if (o.hasDnulls) a.dnulls = (const uint8_t*)(dbase + o.dnulls);
        }
        out[c] = a;
    }
    return PGPU_OK;
}

// ---- coalescing appender --------------------------------------------------------------------------------------------------
// The reference hands operators pages of <= 4096 rows out of RecordPageSource (spi/connector/RecordPageSource.java:84-131) and
// <= 8192-row / ~1 MB batches out of PageProcessor (main/operator/project/PageProcessor.java:56-57). One launch per such page
// would be all launch latency, so host pages are appended, array by array, to a super-page in the current pinned slot
// ("one launch per input page or per coalesced super-page") and the slot is sent — one DMA per array, one
// kernel launch — when it is full, when the page signature changes, or when the operator has to produce output.
inline int32_t rdI32(const char* q) {
    int32_t v;
    memcpy(&v, q, 4);
    return v;
}

int varBytesPerRow(const Type& t) {
    if (t.kind == TypeKind::VARCHAR) return t.precision > 0 ? std::min(t.precision, 64) : 32;
    return 40;  // VARBINARY: intermediate decimal states, 24 / 32 bytes (LongDecimalWithOverflow[AndLong]StateSerializer)
}

const StageLayout& stageLayoutFor(pgpu_pipeline* p, Variant* v, const std::string& key, const std::vector<ColSig>& sig) {
    auto it = p->stageLayouts.find(key);
    if (it != p->stageLayouts.end()) return it->second;
    const GenResult& g = v->gen;
    const Plan& plan = *p->plan;
    size_t rowBytes = 0;
    for (int c : g.usedChannels) {
        const Type& t = plan.input[c];
        if (sig[c].encoding == PGPU_ENC_RLE) continue;
        if (sig[c].encoding == PGPU_ENC_DICTIONARY) rowBytes += 4;
        else rowBytes += t.isVarWidth() ? 4 + (size_t)varBytesPerRow(t) : (size_t)t.width();
        if (sig[c].encoding == PGPU_ENC_PLAIN && sig[c].nullable) rowBytes += 1;
    }
    StageLayout lay;
    // whole tiles of the staged kernel (so that only the last super-page of a stream has a ragged tail), else of the direct one
    const int64_t unit = v->kPageStaged ? g.stagedTile : kTileRows;
    int64_t cap = (int64_t)(p->stageTargetBytes / std::max<size_t>(rowBytes, 1));
    cap = std::max<int64_t>(cap / unit * unit, unit * std::max(p->numSMs, 1));
    lay.capRows = cap;
    lay.cols.assign(g.NC, StageCol{});
    size_t at = 0;
    auto reserve = [&](size_t bytes) {
        const size_t off = at;
        at += align256(std::max<size_t>(bytes, 16));
        return off;
    };
    for (int c : g.usedChannels) {
        const Type& t = plan.input[c];
        StageCol& sc = lay.cols[c];
        sc.encoding = sig[c].encoding;
        sc.nullable = sig[c].nullable;
        if (sig[c].encoding == PGPU_ENC_PLAIN) {
            if (t.isVarWidth()) {
                sc.offsets = reserve((size_t)(cap + 1 + kOffsetsSlack) * 4);
                sc.valuesCap = (size_t)cap * varBytesPerRow(t);
                sc.values = reserve(sc.valuesCap + 32);
            } else {
                sc.width = t.width();
                sc.values = reserve((size_t)cap * t.width());
            }
            if (sig[c].nullable) sc.nulls = reserve((size_t)cap);
        } else {
            if (sig[c].encoding == PGPU_ENC_DICTIONARY) sc.ids = reserve((size_t)cap * 4);
            sc.dvalues = reserve(kDictStageBytes + 32);
            if (t.isVarWidth()) sc.doffsets = reserve(kDictStageBytes + 64);
            else sc.width = t.width();
            if (sig[c].nullable) sc.dnulls = reserve(kDictStageBytes);
        }
    }
    lay.bytes = at;
    return p->stageLayouts.emplace(key, std::move(lay)).first->second;
}

// bytes of the dictionary / RLE value block of a channel, or -1 when it is too large to ride in a super-page
int64_t dictValueBytes(const Type& t, const HostCol& hc) {
    const int64_t dn = hc.dcount;
    if (dn < 0 || (size_t)dn + 1 > kDictStageBytes / 4) return -1;
    if (!t.isVarWidth()) return dn * t.width() <= (int64_t)kDictStageBytes ? dn * t.width() : -1;
    if (dn == 0) return 0;
    const int64_t total = hc.dEndOffsets ? rdI32(hc.doffsets + 4 * (dn - 1)) : (int64_t)rdI32(hc.doffsets + 4 * dn) - rdI32(hc.doffsets);
    return total >= 0 && total <= (int64_t)kDictStageBytes ? total : -1;
}

// start of the dictionary's payload bytes inside dvalues
int64_t dictByteBase(const HostCol& hc) { return hc.dEndOffsets || hc.dcount == 0 || !hc.doffsets ? 0 : rdI32(hc.doffsets); }

bool hostPageCoalescible(pgpu_pipeline* p, Variant* v, const std::vector<ColSig>& sig, const std::vector<HostCol>& hc) {
    const GenResult& g = v->gen;
    if (g.mode == KernelMode::COMPACT) return false;  // filter + project: one output page per input page
    for (int c : g.usedChannels) {
        if (sig[c].encoding == PGPU_ENC_VARCELL) return false;
        if (sig[c].encoding != PGPU_ENC_PLAIN && dictValueBytes(p->plan->input[c], hc[c]) < 0) return false;
    }
    return true;
}

pgpu_status flushPending(pgpu_pipeline* p);

// the dictionaries of `hc` are the ones the pending super-page already carries
bool sameDictionaries(pgpu_pipeline* p, const std::vector<ColSig>& sig, const std::vector<HostCol>& hc) {
    const PendingPage& pd = p->pend;
    const StageSlot& slot = p->slots[p->nextSlot];
    for (int c : pd.v->gen.usedChannels) {
        if (sig[c].encoding == PGPU_ENC_PLAIN) continue;
        const Type& t = p->plan->input[c];
        const StageCol& sc = pd.lay->cols[c];
        const int64_t bytes = dictValueBytes(t, hc[c]);
        if (hc[c].dcount != pd.dcount[c] || bytes != (int64_t)pd.dbytes[c]) return false;
        if (bytes && memcmp((const char*)hc[c].dvalues + dictByteBase(hc[c]), (const char*)slot.host + sc.dvalues, (size_t)bytes)) return false;
        if (hc[c].dnulls && memcmp(hc[c].dnulls, (const char*)slot.host + sc.dnulls, (size_t)hc[c].dcount)) return false;
        if (t.isVarWidth()) {  // same bytes, same boundaries
            const int32_t* have = (const int32_t*)((const char*)slot.host + sc.doffsets);
            const int64_t base = dictByteBase(hc[c]);
            for (int64_t i = 0; i < hc[c].dcount; i++) {
                const int64_t endv = hc[c].dEndOffsets ? rdI32(hc[c].doffsets + 4 * i) : rdI32(hc[c].doffsets + 4 * (i + 1)) - base;
                if (have[i + 1] != (int32_t)endv) return false;
            }
        }
    }
    return true;
}

pgpu_status beginPending(pgpu_pipeline* p, Variant* v, const std::string& key, const std::vector<ColSig>& sig, const std::vector<HostCol>& hc) {
    PendingPage& pd = p->pend;
    const StageLayout& lay = stageLayoutFor(p, v, key, sig);
    StageSlot* slot = nullptr;
    pgpu_status st = acquireSlot(p, &slot);
    if (st) return st;
    st = ensureSlot(p, *slot, lay.bytes, true);
    if (st) return st;
    pd.v = v;
    pd.key = key;
    pd.lay = &lay;
    pd.rows = 0;
    pd.vbytes.assign(v->gen.NC, 0);
    pd.dcount.assign(v->gen.NC, 0);
    pd.dbytes.assign(v->gen.NC, 0);
    // the super-page's dictionaries: copied once, by its first page
    std::vector<CopyTask> tasks;
    char* hbase = (char*)slot->host;
    for (int c : v->gen.usedChannels) {
        if (sig[c].encoding == PGPU_ENC_PLAIN) continue;
        const Type& t = p->plan->input[c];
        const StageCol& sc = lay.cols[c];
        const int64_t bytes = dictValueBytes(t, hc[c]);
        pd.dcount[c] = (int32_t)hc[c].dcount;
        pd.dbytes[c] = (size_t)bytes;
        if (bytes) tasks.push_back(CopyTask{CopyTask::BYTES, hbase + sc.dvalues, (const char*)hc[c].dvalues + dictByteBase(hc[c]), (size_t)bytes, 0});
        if (hc[c].dnulls && hc[c].dcount) tasks.push_back(CopyTask{CopyTask::BYTES, hbase + sc.dnulls, hc[c].dnulls, (size_t)hc[c].dcount, 0});
        if (t.isVarWidth()) {
            int32_t* dst = (int32_t*)(hbase + sc.doffsets);
            dst[0] = 0;
            const int32_t base = (int32_t)dictByteBase(hc[c]);
            if (hc[c].dcount)
                tasks.push_back(CopyTask{CopyTask::OFFSETS32, dst + 1, hc[c].doffsets + (hc[c].dEndOffsets ? 0 : 4), (size_t)hc[c].dcount * 4, -base});
        }
    }
    hostCopy(tasks);
    return PGPU_OK;
}

// Append rows [r0, r0 + n) of a host page to the pending super-page (the caller has made sure they fit).
void appendRows(pgpu_pipeline* p, const std::vector<ColSig>& sig, const std::vector<HostCol>& hc, int64_t r0, int64_t n) {
    PendingPage& pd = p->pend;
    const StageLayout& lay = *pd.lay;
    char* hbase = (char*)p->slots[p->nextSlot].host;
    std::vector<CopyTask> tasks;
    for (int c : pd.v->gen.usedChannels) {
        const Type& t = p->plan->input[c];
        const StageCol& sc = lay.cols[c];
        const HostCol& h = hc[c];
        if (sig[c].encoding == PGPU_ENC_PLAIN) {
            if (t.isVarWidth()) {
                const int32_t b0 = h.endOffsets ? (r0 ? rdI32(h.offsets + 4 * (r0 - 1)) : 0) : rdI32(h.offsets + 4 * r0);
                const char* ends = h.offsets + 4 * (r0 + (h.endOffsets ? 0 : 1));
                const int32_t b1 = rdI32(ends + 4 * (n - 1));
                int32_t* dst = (int32_t*)(hbase + sc.offsets) + pd.rows;
                dst[0] = (int32_t)pd.vbytes[c];
                tasks.push_back(CopyTask{CopyTask::OFFSETS32, dst + 1, ends, (size_t)n * 4, (int32_t)pd.vbytes[c] - b0});
                if (b1 > b0) tasks.push_back(CopyTask{CopyTask::BYTES, hbase + sc.values + pd.vbytes[c], (const char*)h.values + b0, (size_t)(b1 - b0), 0});
                pd.vbytes[c] += b1 - b0;
                p->stats.h2d_bytes += (int64_t)n * 4 + (b1 - b0);
            } else {
                tasks.push_back(CopyTask{CopyTask::BYTES, hbase + sc.values + (size_t)pd.rows * sc.width, (const char*)h.values + (size_t)r0 * sc.width, (size_t)n * sc.width, 0});
                p->stats.h2d_bytes += n * sc.width;
            }
            if (sig[c].nullable) {
                tasks.push_back(CopyTask{CopyTask::BYTES, hbase + sc.nulls + pd.rows, h.nulls + r0, (size_t)n, 0});
                p->stats.h2d_bytes += n;
            }
        } else if (sig[c].encoding == PGPU_ENC_DICTIONARY) {
            tasks.push_back(CopyTask{CopyTask::BYTES, hbase + sc.ids + (size_t)pd.rows * 4, h.ids + (size_t)r0 * 4, (size_t)n * 4, 0});
            p->stats.h2d_bytes += n * 4;
        }
    }
    hostCopy(tasks);
    pd.rows += n;
}

// Send the pending super-page: one DMA per array, one launch of the page kernel.
pgpu_status flushPending(pgpu_pipeline* p) {
    PendingPage& pd = p->pend;
    if (pd.rows == 0) return PGPU_OK;
    CU(cudaSetDevice(p->device));
    Variant* v = pd.v;
    const GenResult& g = v->gen;
    const StageLayout& lay = *pd.lay;
    StageSlot& slot = p->slots[p->nextSlot];
    char* hbase = (char*)slot.host;
    char* dbase = (char*)slot.dev.ptr;
    const int64_t n = pd.rows;
    auto h2d = [&](size_t off, size_t bytes) -> cudaError_t {
        return bytes ? cudaMemcpyAsync(dbase + off, hbase + off, bytes, cudaMemcpyHostToDevice, p->stream) : cudaSuccess;
    };
    DevPage pg;
    pg.rows = n;
    pg.slack = kOffsetsSlack;
    pg.cols.assign(g.NC, ColArg{});
    for (int c : g.usedChannels) {
        const Type& t = p->plan->input[c];
        const StageCol& sc = lay.cols[c];
        ColArg a{};
        if (sc.encoding == PGPU_ENC_PLAIN) {
            if (t.isVarWidth()) {
                int32_t* offs = (int32_t*)(hbase + sc.offsets);
                for (int k = 1; k <= kOffsetsSlack; k++) offs[n + k] = offs[n];  // the staged kernel reads a few offsets past the end
                CU(h2d(sc.offsets, (size_t)(n + 1 + kOffsetsSlack) * 4));
                CU(h2d(sc.values, (size_t)pd.vbytes[c]));
                a.offsets = (const int32_t*)(dbase + sc.offsets);
                a.values = dbase + sc.values;
            } else {
                CU(h2d(sc.values, (size_t)n * sc.width));
                a.values = dbase + sc.values;
            }
            if (sc.nullable) {
                CU(h2d(sc.nulls, (size_t)n));
                a.nulls = (const uint8_t*)(dbase + sc.nulls);
            }
        } else {
            if (sc.encoding == PGPU_ENC_DICTIONARY) {
                CU(h2d(sc.ids, (size_t)n * 4));
                a.ids = (const int32_t*)(dbase + sc.ids);
            }
            CU(h2d(sc.dvalues, pd.dbytes[c]));
            a.dvalues = dbase + sc.dvalues;
            a.dcount = pd.dcount[c];
            if (t.isVarWidth()) {
                CU(h2d(sc.doffsets, (size_t)(pd.dcount[c] + 1) * 4));
                a.doffsets = (const int32_t*)(dbase + sc.doffsets);
            }
            if (sc.nullable) {
                CU(h2d(sc.dnulls, (size_t)pd.dcount[c]));
                a.dnulls = (const uint8_t*)(dbase + sc.dnulls);
            }
        }
        pg.cols[c] = a;
    }
    pd.rows = 0;
    p->stats.super_pages++;
    pgpu_status st = launchPageKernel(p, v, pg);
    if (st) return st;
    return releaseSlot(p, slot);
}

// rows of [r0, rows) that still fit into the pending super-page (row capacity and the payload capacity of var-width channels)
int64_t rowsThatFit(pgpu_pipeline* p, const std::vector<ColSig>& sig, const std::vector<HostCol>& hc, int64_t r0, int64_t rows) {
    const PendingPage& pd = p->pend;
    int64_t n = std::min(rows - r0, pd.lay->capRows - pd.rows);
    for (int c : pd.v->gen.usedChannels) {
        if (n <= 0) break;
        const Type& t = p->plan->input[c];
        if (sig[c].encoding != PGPU_ENC_PLAIN || !t.isVarWidth()) continue;
        const HostCol& h = hc[c];
        const int64_t b0 = h.endOffsets ? (r0 ? rdI32(h.offsets + 4 * (r0 - 1)) : 0) : rdI32(h.offsets + 4 * r0);
        const char* ends = h.offsets + 4 * (r0 + (h.endOffsets ? 0 : 1));
        const int64_t room = (int64_t)pd.lay->cols[c].valuesCap - pd.vbytes[c];
        if ((int64_t)rdI32(ends + 4 * (n - 1)) - b0 <= room) continue;
        int64_t lo = 0, hi = n;  // largest k with bytes(r0, r0 + k) <= room
        while (lo < hi) {
            const int64_t mid = (lo + hi + 1) / 2;
            if ((int64_t)rdI32(ends + 4 * (mid - 1)) - b0 <= room) lo = mid;
            else hi = mid - 1;
        }
        n = lo;
    }
    return std::max<int64_t>(n, 0);
}

pgpu_status appendHostPage(pgpu_pipeline* p, Variant* v, const std::vector<ColSig>& sig, const std::vector<HostCol>& hc, int64_t rows) {
    PendingPage& pd = p->pend;
    const std::string key = sigKey(sig);
    pgpu_status st;
    if (pd.rows > 0 && (pd.key != key || !sameDictionaries(p, sig, hc))) {
        if ((st = flushPending(p))) return st;
    }
    int64_t r0 = 0;
    while (r0 < rows) {
        if (pd.rows == 0 && (st = beginPending(p, v, key, sig, hc))) return st;
        const int64_t n = rowsThatFit(p, sig, hc, r0, rows);
        if (n == 0) {
            if (pd.rows == 0) return fail(p, PGPU_INSUFFICIENT_RESOURCES, "a single var-width value does not fit into a staging slot");
            if ((st = flushPending(p))) return st;
            continue;
        }
        appendRows(p, sig, hc, r0, n);
        r0 += n;
        if (pd.rows >= pd.lay->capRows && (st = flushPending(p))) return st;
    }
    return PGPU_OK;
}

void hostColsOf(const Plan& plan, const pgpu_column* cols, int ncols, std::vector<HostCol>& hc) {
    hc.assign((size_t)ncols, HostCol{});
    for (int c = 0; c < ncols; c++) {
        const pgpu_column& col = cols[c];
        HostCol& h = hc[c];
        h.encoding = col.encoding;
        if (col.encoding == PGPU_ENC_PLAIN) {
            h.values = col.values;
            h.nulls = col.nulls;
            h.offsets = (const char*)col.offsets;
        } else if (col.dictionary) {
            h.ids = (const char*)col.ids;
            h.dvalues = col.dictionary->values;
            h.dnulls = col.dictionary->nulls;
            h.doffsets = (const char*)col.dictionary->offsets;
            h.dcount = col.dictionary->position_count;
        }
    }
}

pgpu_status addPageCommon(pgpu_pipeline* p, const pgpu_column* cols, int32_t ncols, int64_t rows, bool onDevice, uint32_t flags) {
    if (p->device < 0) return fail(p, PGPU_INVALID_STATE, "pipeline was created compile-only (device < 0)");
    if (p->finishing) return fail(p, PGPU_INVALID_STATE, "Operator is already finishing");
    if (!pgpu_needs_input(p)) return fail(p, PGPU_INVALID_STATE, "Operator does not need input");
    std::vector<ColSig> sig;
    pgpu_status st = validatePage(p, cols, ncols, rows, sig);
    if (st) return st;
    CU(cudaSetDevice(p->device));
    p->sawInput = true;
    p->stats.input_pages++;
    p->stats.input_rows += rows;
    if (rows == 0) return PGPU_OK;
    Variant* v = nullptr;
    st = getVariant(p, sig, &v);
    if (st) return st;
    if (onDevice) {
        for (int c : v->gen.usedChannels) {
            const pgpu_column& col = cols[c];
            auto mis = [](const void* q, size_t a) { return q && ((uintptr_t)q % a) != 0; };
            if (mis(col.values, 16) || mis(col.ids, 16) || mis(col.offsets, 16) || mis(col.nulls, 4))
                return fail(p, PGPU_INVALID_ARGUMENT, "device blocks must be 16-byte aligned (nulls: 4-byte)");
        }
        if ((st = flushPending(p))) return st;
        return processDevicePage(p, v, cols, rows);
    }
    for (int c : v->gen.usedChannels)
        if (cols[c].encoding == PGPU_ENC_VARCELL) return fail(p, PGPU_NOT_SUPPORTED, "not supported on the GPU path: 16-byte cell columns in host pages (device-to-device encoding)");
    const bool pinned = (flags & PGPU_PAGE_PINNED) != 0;
    if (!pinned) {
        // ordinary (pageable) memory: append to the super-page in the pinned ring
        std::vector<HostCol> hc;
        hostColsOf(*p->plan, cols, ncols, hc);
        if (hostPageCoalescible(p, v, sig, hc)) return appendHostPage(p, v, sig, hc, rows);
    }
    if ((st = flushPending(p))) return st;
    // caller-pinned pages (DMA straight from the caller's buffers), filter + project pages, oversized dictionaries: staged in
    // chunks through the ring, one launch per chunk
    const bool compact = v->gen.mode == KernelMode::COMPACT;
    int64_t chunk = compact ? rows : kHostChunkRows;
    for (int64_t r0 = 0; r0 < rows; r0 += chunk) {
        int64_t n = std::min(chunk, rows - r0);
        StageSlot* slotp = nullptr;
        if ((st = acquireSlot(p, &slotp))) return st;
        StageSlot& slot = *slotp;
        DevPage pg;
        pg.rows = n;
        st = stageChunk(p, v, cols, r0, n, pinned, slot, pg.cols);
        if (st) return st;
        if (compact) {
            st = runCompact(p, v, pg);
        } else {
            st = launchPageKernel(p, v, pg);
        }
        if (st) return st;
        if ((st = releaseSlot(p, slot))) return st;
    }
    return PGPU_OK;
}

// any compiled variant whose kernels match the table's slot format (emit / rehash / count kernels do not depend on the page signature)
Variant* tableVariant(pgpu_pipeline* p) {
    const bool buckets = p->tableMode == pgpu_pipeline::TABLE_BUCKETS;
    for (auto& kv : p->variants)
        if (kv.second.gen.buckets == buckets) return &kv.second;
    return nullptr;
}

// PARTIAL step: once the row-count upper bound crosses the flush threshold, learn the real group count
pgpu_status updatePartialFill(pgpu_pipeline* p) {
    const Plan& plan = *p->plan;
    if (plan.hasAgg && outputIsIntermediate(plan.step) && plan.maxPartialGroups > 0 && !plan.groupChannels.empty() &&
        p->groupsKnown + p->rowsSinceKnown + p->pend.rows >= plan.maxPartialGroups) {
        pgpu_status st = flushPending(p);  // rows still waiting in the staging slot count: aggregate them, then look
        if (st) return st;
        if (p->dSlots) return refreshGroupCount(p, tableVariant(p), true);
    }
    return PGPU_OK;
}

// materialise the aggregation result into the device output buffers
pgpu_status buildAggOutput(pgpu_pipeline* p) {
    const GenResult& g = p->layout;
    CU(cudaSetDevice(p->device));
    {
        pgpu_status st = flushPending(p);
        if (st) return st;
    }
    // any compiled variant has the emit kernels (layout is signature independent)
    Variant* v = tableVariant(p);
    if (!v) {
        std::vector<ColSig> sig(p->plan->input.size());
        pgpu_status st = getVariant(p, sig, &v);
        if (st) return st;
    }
    CU(cudaMemsetAsync(p->dOutCount, 0, 8, p->stream));
    if (g.mode == KernelMode::AGG_GLOBAL) {
        pgpu_status st = ensureOutBuffers(p, 1);
        if (st) return st;
        std::vector<unsigned char> oa = outArgsBlock(p);
        Scratch sc = scratchOf(p);
        void* args[] = {&sc, oa.data()};
        st = launch(p, v->kEmit, 1, 32, 0, args);
        if (st) return st;
        p->outRowsTotal = 1;
    } else {
        if (p->streamMode == pgpu_pipeline::STREAM_ON && p->runRows > 0 && !p->dSlots) {
            // streaming runs: neighbours with the same key merge, every group is finalised straight out of the buffer
            v = streamVariant(p);
            if (!v) return fail(p, PGPU_CUDA_ERROR, "internal: buffered runs without an emit kernel");
            pgpu_status st = ensureOutBuffers(p, p->runRows + p->extraOutRows);
            if (st) return st;
            CU(cudaMemsetAsync(p->dViolations, 0, 4, p->stream));
            std::vector<unsigned char> oa = outArgsBlock(p);
            Scratch sc = scratchOf(p);
            const uint64_t* buf = (const uint64_t*)p->runBuf.ptr;
            int64_t n = p->runRows;
            void* args[] = {&buf, &n, oa.data(), &sc, &p->dViolations};
            int grid = (int)std::min<int64_t>((n + kThreads * 4 - 1) / (kThreads * 4), (int64_t)p->numSMs * 8);  // 4 = kRunsEmitPer
            st = launch(p, v->kRunsEmit, grid, kThreads, 0, args);
            if (st) return st;
            CU(cudaMemcpyAsync(&p->hostScalars[0], p->dOutCount, 8, cudaMemcpyDeviceToHost, p->stream));
            CU(cudaMemcpyAsync(&p->hostScalars[14], p->dViolations, 4, cudaMemcpyDeviceToHost, p->stream));
            CU(cudaStreamSynchronize(p->stream));
            if ((uint32_t)p->hostScalars[14] == 0) {
                p->outRowsTotal = p->groupsKnown = (int64_t)p->hostScalars[0];
                st = checkDeviceErrors(p);
                if (st) return st;
                p->outputReady = true;
                p->outRowsServed = 0;
                return PGPU_OK;
            }
            // the key order decreases somewhere: a key may occur in two places. Everything goes through the table after all.
            if ((st = streamToTable(p, v))) return st;
            CU(cudaMemsetAsync(p->dOutCount, 0, 8, p->stream));
        }
        if (!p->dSlots) {
            p->outRowsTotal = 0;
        } else {
            // a bucketed table knows its groups only by counting presence bits. When the probe saw well-filled buckets the
            // upper bound (8 per claimed bucket) is close enough to size the output with; the emit kernel reports the real count
            const bool bounded = p->tableMode == pgpu_pipeline::TABLE_BUCKETS && p->bucketOccupancy >= 4.0;
            pgpu_status st = refreshGroupCount(p, v, !bounded);
            if (st) return st;
            CU(cudaMemsetAsync(p->dOutCount, 0, 8, p->stream));  // the group count of a bucketed table went through it
            st = ensureOutBuffers(p, p->groupsKnown + p->extraOutRows);  // extraOutRows: rows the exchange appends behind the table's groups
            if (st) return st;
            std::vector<unsigned char> oa = outArgsBlock(p);
            Scratch sc = scratchOf(p);
            TableArgs t = tableOf(p);
            int64_t first = 0, nslots = (int64_t)p->capacity;
            void* args[] = {&t, oa.data(), &sc, &first, &nslots};
            int grid = (int)std::min<uint64_t>((p->capacity + kThreads - 1) / kThreads, (uint64_t)p->numSMs * 8);
            st = launch(p, v->kEmit, grid, kThreads, 0, args);
            if (st) return st;
            p->outRowsTotal = p->groupsKnown;
            if (bounded) {
                CU(cudaMemcpyAsync(&p->hostScalars[0], p->dOutCount, 8, cudaMemcpyDeviceToHost, p->stream));
                p->outRowsTotal = -1;  // read below
            }
        }
    }
    CU(cudaStreamSynchronize(p->stream));
    if (p->outRowsTotal < 0) p->outRowsTotal = p->groupsKnown = (int64_t)p->hostScalars[0];
    pgpu_status st = checkDeviceErrors(p);
    if (st) return st;
    p->outputReady = true;
    p->outRowsServed = 0;
    return PGPU_OK;
}

void resetAggState(pgpu_pipeline* p) {
    if (p->device < 0) return;
    cudaSetDevice(p->device);
    if (p->dState) cudaMemsetAsync(p->dState, 0, 8 * 64, p->stream);
    if (p->dCounters) cudaMemsetAsync(p->dCounters, 0, 16, p->stream);
    if (p->dSlots) cudaMemsetAsync(p->dSlots, 0, (size_t)p->capacity * p->layout.SLOTW * 8, p->stream);
    if (p->dErr) cudaMemsetAsync(p->dErr, 0, 4, p->stream);
    if (p->dTicket) cudaMemsetAsync(p->dTicket, 0, 4, p->stream);
    p->groupsKnown = 0;
    p->slotsKnown = 0;
    p->rowsSinceKnown = 0;
}

bool isPartialFlushDue(const pgpu_pipeline* p) {
    return p->plan->hasAgg && outputIsIntermediate(p->plan->step) && p->plan->maxPartialGroups > 0 && !p->plan->groupChannels.empty() &&
           (p->groupsKnown + p->rowsSinceKnown + p->pend.rows) >= p->plan->maxPartialGroups;
}

}  // namespace

// for translation units that have no pipeline handle to hang an error text on (serde.cu: pgpu_expand_nulls_static)
namespace pgpu {
void setLastGlobalError(const std::string& m) { g_lastGlobalError = m; }
}  // namespace pgpu

// ================================================================================================ C ABI
extern "C" {

const char* pgpu_version(void) { return "prestogpu 0.1 (sm_100a)"; }

int32_t pgpu_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

const char* pgpu_last_global_error(void) { return g_lastGlobalError.c_str(); }

pgpu_status pgpu_pipeline_create(const uint8_t* plan, size_t plan_len, int32_t device, pgpu_pipeline** out) {
    if (!plan || !out) {
        g_lastGlobalError = "null argument";
        return PGPU_INVALID_ARGUMENT;
    }
    *out = nullptr;
    auto p = std::make_unique<pgpu_pipeline>();
    try {
        p->plan = parsePlan(std::string((const char*)plan, plan_len));
    } catch (const PlanError& e) {
        g_lastGlobalError = e.what();
        return (pgpu_status)e.status;
    }
    p->device = device;
    if (const char* e = getenv("PGPU_STAGE_MB")) p->stageTargetBytes = (size_t)std::max(1, atoi(e)) << 20;
    pgpu_pipeline* pp = p.get();
    auto cuFail = [&](cudaError_t e, const char* what) {
        g_lastGlobalError = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what;
        return PGPU_CUDA_ERROR;
    };
    if (device >= 0) {
        cudaError_t e;
        if ((e = cudaSetDevice(device)) != cudaSuccess) return cuFail(e, "cudaSetDevice");
        cudaDeviceProp prop;
        if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return cuFail(e, "cudaGetDeviceProperties");
        if (prop.major != 10) {
            g_lastGlobalError = std::string("prestogpu kernels are built for sm_100a only; device is sm_") + std::to_string(prop.major) + std::to_string(prop.minor);
            return PGPU_NOT_SUPPORTED;
        }
        p->numSMs = prop.multiProcessorCount;
        if ((e = cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking)) != cudaSuccess) return cuFail(e, "cudaStreamCreate");
        if ((e = cudaEventCreate(&p->evA)) != cudaSuccess) return cuFail(e, "cudaEventCreate");
        if ((e = cudaEventCreate(&p->evB)) != cudaSuccess) return cuFail(e, "cudaEventCreate");
        void* scratch = nullptr;
        // err(4) ticket(4) out_count(8) counters(16) state(64*8)
        if ((e = cudaMalloc(&scratch, 1024)) != cudaSuccess) return cuFail(e, "cudaMalloc");
        if ((e = cudaMemset(scratch, 0, 1024)) != cudaSuccess) return cuFail(e, "cudaMemset");
        char* s = (char*)scratch;
        p->dErr = (uint32_t*)s;
        p->dTicket = (uint32_t*)(s + 8);
        p->dOutCount = (uint64_t*)(s + 16);
        p->dCounters = (uint64_t*)(s + 32);
        p->dState = (uint64_t*)(s + 64);
        if ((e = cudaMallocHost((void**)&p->hostScalars, 256)) != cudaSuccess) return cuFail(e, "cudaMallocHost");
        memset(p->hostScalars, 0, 256);
    }
    // compile the default variant now: plan-time rejection (NOT_SUPPORTED / COMPILER_ERROR) happens here
    std::vector<ColSig> sig(p->plan->input.size());
    Variant* v = nullptr;
    pgpu_status st = getVariant(pp, sig, &v);
    if (st) {
        g_lastGlobalError = p->err;
        pgpu_pipeline_destroy(p.release());
        return st;
    }
    if (!v->gen.bucketsPossible) {
        p->tableMode = pgpu_pipeline::TABLE_PER_KEY;
    } else if (const char* force = getenv("PGPU_TUNE_BUCKETS")) {
        if (!strcmp(force, "auto")) {
            // decided by the first page's keys
        } else if (*force && atoi(force) != 0) {  // bucketed table from the start (also what build() uses to warm the kernel cache)
            p->tableMode = pgpu_pipeline::TABLE_BUCKETS;
            p->stats.table_bucketed = 1;
            if (const char* off = getenv("PGPU_TUNE_BUCKET_OFF")) p->bucketOff = (uint64_t)(atoi(off) & 7);
            p->bucketOccupancy = 4.0;  // no probe: a middling guess, the table grows when it is wrong
            p->haveLayout = false;
            st = getVariant(pp, sig, &v);
            if (st) {
                g_lastGlobalError = p->err;
                pgpu_pipeline_destroy(p.release());
                return st;
            }
        } else if (*force) {
            p->tableMode = pgpu_pipeline::TABLE_PER_KEY;
        }
    }
    if (p->layout.NW > 60) {
        g_lastGlobalError = "not supported on the GPU path: more than 60 accumulator words";
        pgpu_pipeline_destroy(p.release());
        return PGPU_NOT_SUPPORTED;
    }
    *out = p.release();
    return PGPU_OK;
}

void pgpu_pipeline_destroy(pgpu_pipeline* p) {
    if (!p) return;
    if (p->singleTwin) pgpu_pipeline_destroy(p->singleTwin);
    if (p->device >= 0) {
        cudaSetDevice(p->device);
        if (p->stream) cudaStreamSynchronize(p->stream);
        for (auto& kv : p->variants)
            if (kv.second.lib) cudaLibraryUnload(kv.second.lib);
        if (p->dErr) cudaFree(p->dErr);
        if (p->dPartials) cudaFree(p->dPartials);
        if (p->tileStatus.ptr) cudaFree(p->tileStatus.ptr);
        if (p->failLog.ptr) cudaFree(p->failLog.ptr);
        if (p->sweepBuf.ptr) cudaFree(p->sweepBuf.ptr);
        if (p->runBuf.ptr) cudaFree(p->runBuf.ptr);
        if (p->runCounts.ptr) cudaFree(p->runCounts.ptr);
        if (p->runOffsets.ptr) cudaFree(p->runOffsets.ptr);
        if (p->dViolations) cudaFree(p->dViolations);
        if (p->sweepMeta.ptr) cudaFree(p->sweepMeta.ptr);
        if (p->dSlots) cudaFree(p->dSlots);
        if (p->hostScalars) cudaFreeHost(p->hostScalars);
        for (auto& s : p->slots) {
            if (s.dev.ptr) cudaFree(s.dev.ptr);
            if (s.host) cudaFreeHost(s.host);
            if (s.done) cudaEventDestroy(s.done);
        }
        if (p->outArena.ptr) cudaFree(p->outArena.ptr);
        if (p->outHostArena.ptr) cudaFreeHost(p->outHostArena.ptr);
        if (p->serdeArena.ptr) cudaFree(p->serdeArena.ptr);
        if (p->serdeOutArena.ptr) cudaFree(p->serdeOutArena.ptr);
        if (p->parquetArena.ptr) cudaFree(p->parquetArena.ptr);
        if (p->serdeHost.ptr) cudaFreeHost(p->serdeHost.ptr);
        if (p->serdeTotals.ptr) cudaFreeHost(p->serdeTotals.ptr);
        if (p->partArena.ptr) cudaFree(p->partArena.ptr);
        for (void* q : p->partRetired) cudaFree(q);
        for (auto* b : {&p->partPid, &p->partDst, &p->partCounts})
            if (b->ptr) cudaFree(b->ptr);
        if (p->hostPart) cudaFreeHost(p->hostPart);
        for (auto* vec : {&p->outHost, &p->outHostNulls, &p->outHostAux})
            for (auto& b : *vec)
                if (b.ptr) cudaFreeHost(b.ptr);
        if (p->evA) cudaEventDestroy(p->evA);
        if (p->evB) cudaEventDestroy(p->evB);
        if (p->stream) cudaStreamDestroy(p->stream);
    }
    delete p;
}

const char* pgpu_last_error(const pgpu_pipeline* p) { return p ? p->err.c_str() : "null pipeline"; }

int32_t pgpu_needs_input(const pgpu_pipeline* p) {
    if (!p || p->finishing || p->finished) return 0;
    if (p->outputReady) return 0;  // HashAggregationOperator.needsInput: outputPages != null => false
    if (isPartialFlushDue(p)) return 0;  // aggregationBuilder.isFull()
    return 1;
}

pgpu_status pgpu_add_page(pgpu_pipeline* p, const pgpu_column* cols, int32_t ncols, int64_t rows, uint32_t flags) {
    if (!p) return PGPU_INVALID_ARGUMENT;
    pgpu_status st = addPageCommon(p, cols, ncols, rows, false, flags);
    return st ? st : updatePartialFill(p);
}

pgpu_status pgpu_add_device_page(pgpu_pipeline* p, const pgpu_column* cols, int32_t ncols, int64_t rows) {
    if (!p) return PGPU_INVALID_ARGUMENT;
    pgpu_status st = addPageCommon(p, cols, ncols, rows, true, 0);
    return st ? st : updatePartialFill(p);
}

pgpu_status pgpu_reserve_input(pgpu_pipeline* p, const int32_t* encodings, const uint8_t* nullable, int32_t ncols, int64_t rows) {
    if (!p || !encodings || !nullable) return PGPU_INVALID_ARGUMENT;
    if (ncols != (int)p->plan->input.size()) return fail(p, PGPU_INVALID_ARGUMENT, "signature length differs from the plan's input channels");
    std::vector<ColSig> sig(ncols);
    for (int c = 0; c < ncols; c++) {
        sig[c].encoding = encodings[c];
        sig[c].nullable = nullable[c] != 0;
    }
    Variant* v = nullptr;
    pgpu_status st = getVariant(p, sig, &v);  // compiles on first use
    if (st || p->device < 0) return st;
    CU(cudaSetDevice(p->device));
    PendingPage& pd = p->pend;
    if (pd.rows > 0 && (pd.key != sigKey(sig) || pd.rows + rows > pd.lay->capRows)) {
        if ((st = flushPending(p))) return st;
    }
    if (pd.rows == 0) {  // the next page starts a super-page: its slot must be free
        StageSlot* slot = nullptr;
        if ((st = acquireSlot(p, &slot))) return st;
    }
    return PGPU_OK;
}

int32_t pgpu_is_blocked(const pgpu_pipeline* p) {
    if (!p || p->device < 0 || !p->stream) return 0;
    if (cudaSetDevice(p->device) != cudaSuccess) return 0;
    const cudaError_t e = cudaStreamQuery(p->stream);
    if (e == cudaErrorNotReady) return 1;
    if (e != cudaSuccess) cudaGetLastError();
    return 0;
}

namespace {
struct UnblockedCall {
    pgpu_callback fn;
    void* user;
};
void CUDART_CB unblockedTrampoline(void* arg) {
    UnblockedCall* c = (UnblockedCall*)arg;
    c->fn(c->user);
    delete c;
}
}  // namespace

pgpu_status pgpu_on_unblocked(pgpu_pipeline* p, pgpu_callback fn, void* user) {
    if (!p || !fn) return PGPU_INVALID_ARGUMENT;
    if (p->device < 0) {
        fn(user);
        return PGPU_OK;
    }
    CU(cudaSetDevice(p->device));
    UnblockedCall* c = new UnblockedCall{fn, user};
    cudaError_t e = cudaLaunchHostFunc(p->stream, unblockedTrampoline, c);
    if (e != cudaSuccess) {
        delete c;
        p->err = std::string("CUDA error: ") + cudaGetErrorString(e) + " at cudaLaunchHostFunc";
        return PGPU_CUDA_ERROR;
    }
    return PGPU_OK;
}

// ---- pages in the reference's wire format -----------------------------------------------------------------------------------
namespace {

// One block of a serialized page: where its pieces lie in the payload (byte offsets). Parsed on the host from the few
// header bytes only; the arrays themselves are never touched by the CPU.
struct WireBlock {
    enum Kind { FLAT, VARWIDTH, DICTIONARY, RLE } kind = FLAT;
    int width = 0;           // FLAT: bytes per value
    int64_t n = 0;           // positionCount
    bool hasNulls = false;
    size_t bitsOff = 0;      // packed null flags, (n + 7) / 8 bytes
    int64_t nonNull = 0;     // FLAT with nulls: values present
    size_t valuesOff = 0;    // FLAT: values; VARWIDTH: bytes; DICTIONARY: ids
    size_t offsetsOff = 0;   // VARWIDTH: n int32 END offsets
    int64_t dataLen = 0;     // VARWIDTH: total bytes
    std::unique_ptr<WireBlock> child;  // DICTIONARY: dictionary; RLE: value
};

struct WireReader {
    const uint8_t* d;
    size_t len, pos = 0;
    bool ok = true;
    int32_t i32() {
        if (pos + 4 > len) { ok = false; return 0; }
        int32_t v;
        memcpy(&v, d + pos, 4);
        pos += 4;
        return v;
    }
    size_t skip(size_t n) {
        size_t at = pos;
        if (n > len - std::min(pos, len)) { ok = false; return at; }
        pos += n;
        return at;
    }
};

// main/metadata/InternalBlockEncodingSerde.java:43-53 + spi/block/*BlockEncoding.readBlock
bool parseWireBlock(WireReader& r, WireBlock& b, int depth, std::string& why) {
    const int32_t nameLen = r.i32();
    if (!r.ok || nameLen < 0 || nameLen > 32) { why = "corrupt block encoding name"; return false; }
    const size_t nameAt = r.skip((size_t)nameLen);
    if (!r.ok) { why = "truncated page"; return false; }
    const std::string name((const char*)r.d + nameAt, (size_t)nameLen);
    b.n = r.i32();
    if (!r.ok || b.n < 0) { why = "truncated page"; return false; }
    auto nulls = [&]() {  // spi/block/EncoderUtil.java:33-70
        const size_t at = r.skip(1);
        if (!r.ok) return;
        b.hasNulls = r.d[at] != 0;
        if (b.hasNulls) b.bitsOff = r.skip((size_t)((b.n + 7) / 8));
    };
    int width = name == "LONG_ARRAY" ? 8 : name == "INT_ARRAY" ? 4 : name == "BYTE_ARRAY" ? 1 : name == "INT128_ARRAY" ? 16 : 0;
    if (width) {
        b.kind = WireBlock::FLAT;
        b.width = width;
        nulls();
        b.nonNull = b.n;
        if (r.ok && b.hasNulls) b.nonNull = r.i32();
        if (!r.ok || b.nonNull < 0 || b.nonNull > b.n) { why = "corrupt " + name + " block"; return false; }
        b.valuesOff = r.skip((size_t)b.nonNull * (size_t)width);
        // The device scatters the nonNull values to the positions whose null bit is clear: the two must agree, or the kernel would
        // read past the values (the reference fails in readBlock with an index error instead). n / 8 bytes of popcount.
        if (r.ok && b.hasNulls) {
            int64_t set = 0;
            const uint8_t* bits = r.d + b.bitsOff;
            for (int64_t k = 0; k < b.n / 8; k++) set += __builtin_popcount(bits[k]);
            if (b.n % 8) set += __builtin_popcount(bits[b.n / 8] >> (8 - b.n % 8));  // most significant bit first (EncoderUtil.encodeNullsAsBits)
            if (set != b.n - b.nonNull) { why = "corrupt " + name + " block: " + std::to_string(b.nonNull) + " values for " + std::to_string(b.n - set) + " non-null positions"; return false; }
        }
    } else if (name == "VARIABLE_WIDTH") {
        b.kind = WireBlock::VARWIDTH;
        b.offsetsOff = r.skip((size_t)b.n * 4);
        nulls();
        b.dataLen = r.i32();
        if (!r.ok || b.dataLen < 0) { why = "corrupt VARIABLE_WIDTH block"; return false; }
        b.valuesOff = r.skip((size_t)b.dataLen);
        if (r.ok && b.n > 0) {  // end offsets: ascending and inside the data (VariableWidthBlockEncoding.readBlock trusts them; the device would follow them)
            int32_t bad = 0, last;
            memcpy(&last, r.d + b.offsetsOff + (size_t)(b.n - 1) * 4, 4);
            hostCopy({CopyTask{CopyTask::CHECK_ASCENDING32, &bad, r.d + b.offsetsOff, (size_t)b.n * 4, 0}});  // on the copy pool's threads: 4 bytes per row
            if (bad || last > b.dataLen) { why = "corrupt VARIABLE_WIDTH block: offsets are not ascending or exceed the data length"; return false; }
        }
    } else if (name == "DICTIONARY" || name == "RLE") {
        if (depth > 0) { why = "not supported on the GPU path: nested dictionary/RLE encodings"; return false; }
        b.kind = name == "RLE" ? WireBlock::RLE : WireBlock::DICTIONARY;
        b.child.reset(new WireBlock());
        if (!parseWireBlock(r, *b.child, depth + 1, why)) return false;
        if (b.kind == WireBlock::DICTIONARY) {
            b.valuesOff = r.skip((size_t)b.n * 4);
            r.skip(24);  // DictionaryId: most / least significant bits, sequence id
            if (r.ok && b.n > 0) {  // ids index the dictionary on the device (and its shared-memory copy): every id must be inside it
                int32_t bad = 0;
                hostCopy({CopyTask{CopyTask::CHECK_IDS32, &bad, r.d + b.valuesOff, (size_t)b.n * 4, (int32_t)std::min<int64_t>(b.child->n, INT32_MAX)}});
                if (bad) { why = "corrupt DICTIONARY block: an id lies outside the dictionary of " + std::to_string(b.child->n) + " entries"; return false; }
            }
        } else if (b.child->n != 1) {
            why = "RLE value block must have exactly one position";
            return false;
        }
    } else {
        why = "not supported on the GPU path: block encoding " + name;
        return false;
    }
    if (!r.ok) { why = "truncated page"; return false; }
    return true;
}

}  // namespace

pgpu_status pgpu_add_serialized_page_marked(pgpu_pipeline* p, const uint8_t* data, size_t len, int64_t position_count, int32_t markers, int64_t uncompressed_size) {
    if (!p || (!data && len)) return PGPU_INVALID_ARGUMENT;
    if (markers & ~(PGPU_PAGE_MARKER_COMPRESSED | PGPU_PAGE_MARKER_ENCRYPTED)) return fail(p, PGPU_INVALID_ARGUMENT, "unknown page codec markers");
    if (markers & PGPU_PAGE_MARKER_ENCRYPTED) return fail(p, PGPU_NOT_SUPPORTED, "not supported on the GPU path: encrypted pages (the spill cipher is the caller's)");
    if (!(markers & PGPU_PAGE_MARKER_COMPRESSED)) return pgpu_add_serialized_page(p, data, len, position_count);
    // PagesSerde.deserialize :156-165: decompress into a buffer of getUncompressedSizeInBytes(), the sizes must agree
    if (uncompressed_size < 0 || uncompressed_size > ((int64_t)1 << 31)) return fail(p, PGPU_INVALID_ARGUMENT, "uncompressed size of a compressed page");
    std::vector<uint8_t> raw((size_t)uncompressed_size);
    if (!pq::lz4Decompress(data, len, raw.data(), raw.size())) return fail(p, PGPU_INVALID_ARGUMENT, "corrupt page: the LZ4 slice does not decompress to its declared size");
    return pgpu_add_serialized_page(p, raw.data(), raw.size(), position_count);
}

pgpu_status pgpu_add_serialized_page(pgpu_pipeline* p, const uint8_t* data, size_t len, int64_t position_count) {
    if (!p || (!data && len)) return PGPU_INVALID_ARGUMENT;
    if (p->device < 0) return fail(p, PGPU_INVALID_STATE, "pipeline was created compile-only (device < 0)");
    if (p->finishing) return fail(p, PGPU_INVALID_STATE, "Operator is already finishing");
    if (!pgpu_needs_input(p)) return fail(p, PGPU_INVALID_STATE, "Operator does not need input");
    const Plan& plan = *p->plan;
    WireReader r{data, len};
    const int32_t nblocks = r.i32();  // PagesSerdeUtil.readRawPage
    if (!r.ok || nblocks != (int)plan.input.size())
        return fail(p, PGPU_INVALID_ARGUMENT, "page has " + std::to_string(nblocks) + " channels, plan expects " + std::to_string(plan.input.size()));
    std::vector<WireBlock> blocks((size_t)nblocks);
    for (int c = 0; c < nblocks; c++) {
        std::string why;
        if (!parseWireBlock(r, blocks[c], 0, why))
            return fail(p, why.rfind("not supported", 0) == 0 ? PGPU_NOT_SUPPORTED : PGPU_INVALID_ARGUMENT, why + " (channel " + std::to_string(c) + ")");
        if (blocks[c].n != position_count) return fail(p, PGPU_INVALID_ARGUMENT, "channel " + std::to_string(c) + " position count differs from the page");
        const WireBlock& leaf = blocks[c].child ? *blocks[c].child : blocks[c];
        const Type& t = plan.input[c];
        const bool match = t.isVarWidth() ? leaf.kind == WireBlock::VARWIDTH : (leaf.kind == WireBlock::FLAT && leaf.width == t.width());
        if (!match) return fail(p, PGPU_INVALID_ARGUMENT, "channel " + std::to_string(c) + " block encoding does not match the plan type (" + t.str() + ")");
    }
    if (r.pos != len) return fail(p, PGPU_INVALID_ARGUMENT, "trailing bytes after the last block");
    CU(cudaSetDevice(p->device));
    // Blocks without null flags carry their arrays as they are (LongArrayBlockEncoding.java:39-60: values of every position;
    // VariableWidthBlockEncoding.java:37-58: end offsets + bytes): such a page is appended to the current super-page straight
    // out of the payload, like a page of host blocks — no per-page DMA, no per-page decode launches.
    bool plainArrays = true;
    for (int c = 0; c < nblocks; c++) plainArrays &= !blocks[c].hasNulls && !(blocks[c].child && blocks[c].child->hasNulls);
    if (plainArrays) {
        std::vector<ColSig> sig((size_t)nblocks);
        std::vector<HostCol> hc((size_t)nblocks);
        const char* base = (const char*)data;
        bool sane = true;
        for (int c = 0; c < nblocks; c++) {
            const WireBlock& b = blocks[c];
            HostCol& h = hc[c];
            sig[c].encoding = b.kind == WireBlock::DICTIONARY ? PGPU_ENC_DICTIONARY : b.kind == WireBlock::RLE ? PGPU_ENC_RLE : PGPU_ENC_PLAIN;
            sig[c].nullable = false;
            h.encoding = sig[c].encoding;
            auto lastEnd = [&](const WireBlock& w) { return w.n ? rdI32(base + w.offsetsOff + 4 * (size_t)(w.n - 1)) : 0; };
            if (!b.child) {
                h.values = base + b.valuesOff;
                if (b.kind == WireBlock::VARWIDTH) {
                    h.offsets = base + b.offsetsOff;
                    h.endOffsets = true;
                    sane &= lastEnd(b) <= b.dataLen;
                }
            } else {
                const WireBlock& d = *b.child;
                h.ids = base + b.valuesOff;
                h.dvalues = base + d.valuesOff;
                h.dcount = d.n;
                if (d.kind == WireBlock::VARWIDTH) {
                    h.doffsets = base + d.offsetsOff;
                    h.dEndOffsets = true;
                    sane &= lastEnd(d) <= d.dataLen;
                }
            }
        }
        if (!sane) return fail(p, PGPU_INVALID_ARGUMENT, "corrupt VARIABLE_WIDTH block: offsets exceed the data length");
        Variant* v = nullptr;
        pgpu_status st = getVariant(p, sig, &v);
        if (st) return st;
        if (position_count > 0 && hostPageCoalescible(p, v, sig, hc)) {
            p->sawInput = true;
            p->stats.input_pages++;
            p->stats.input_rows += position_count;
            st = appendHostPage(p, v, sig, hc, position_count);
            return st ? st : updatePartialFill(p);
        }
    }
    // device arena: [payload][decoded arrays, each 256-byte aligned]
    size_t need = align256(len);
    auto reserve = [&](size_t bytes) {
        size_t at = need;
        need += align256(std::max<size_t>(bytes, 16));
        return at;
    };
    struct Placed {
        size_t values = 0, nulls = 0, offsets = 0, compact = 0, scratch = 0;
    };
    auto place = [&](const WireBlock& b) {
        Placed q;
        if (b.kind == WireBlock::FLAT) {
            q.values = reserve((size_t)b.n * b.width);
            if (b.hasNulls) {
                q.nulls = reserve((size_t)b.n);
                q.compact = reserve((size_t)b.nonNull * b.width);
                q.scratch = reserve(serdeScratchWords(b.n) * 4);
            }
        } else if (b.kind == WireBlock::VARWIDTH) {
            q.offsets = reserve(((size_t)b.n + 1) * 4 + 16);
            q.values = reserve((size_t)b.dataLen + 16);
            if (b.hasNulls) q.nulls = reserve((size_t)b.n);
        } else if (b.kind == WireBlock::DICTIONARY) {
            q.values = reserve((size_t)b.n * 4);
        }
        return q;
    };
    std::vector<Placed> at((size_t)nblocks), atChild((size_t)nblocks);
    for (int c = 0; c < nblocks; c++) {
        at[c] = place(blocks[c]);
        if (blocks[c].child) atChild[c] = place(*blocks[c].child);
    }
    pgpu_status st = ensureDev(p, p->serdeArena, need);
    if (st) return st;
    char* base = (char*)p->serdeArena.ptr;
    if (len) CU(cudaMemcpyAsync(base, data, len, cudaMemcpyHostToDevice, p->stream));  // the whole page in one DMA
    p->stats.h2d_bytes += (int64_t)len;
    // decode on the device: aligned copies of the arrays, null flags unpacked, sparse values re-expanded
    auto d2d = [&](size_t dst, size_t src, size_t bytes) -> cudaError_t {
        return bytes ? cudaMemcpyAsync(base + dst, base + src, bytes, cudaMemcpyDeviceToDevice, p->stream) : cudaSuccess;
    };
    auto decode = [&](const WireBlock& b, const Placed& q, const Type& t, pgpu_column& col) -> cudaError_t {
        cudaError_t e = cudaSuccess;
        memset(&col, 0, sizeof col);
        col.encoding = PGPU_ENC_PLAIN;
        col.type_kind = (int32_t)t.kind;
        col.precision = t.precision;
        col.scale = t.scale;
        col.position_count = b.n;
        if (b.kind == WireBlock::FLAT) {
            col.values = base + q.values;
            if (!b.hasNulls) return d2d(q.values, b.valuesOff, (size_t)b.n * b.width);
            col.nulls = (const uint8_t*)(base + q.nulls);
            if ((e = d2d(q.compact, b.valuesOff, (size_t)b.nonNull * b.width))) return e;
            p->stats.kernel_launches += 3;
            return expandNulls(p->stream, (const uint8_t*)(base + b.bitsOff), b.n, base + q.compact, base + q.values, (uint8_t*)(base + q.nulls), b.width, (uint32_t*)(base + q.scratch));
        }
        // VARWIDTH: offsets[0] = 0, offsets[1..n] = the END offsets of the wire (VariableWidthBlockEncoding.readBlock)
        col.offsets = (const int32_t*)(base + q.offsets);
        col.values = base + q.values;
        if ((e = cudaMemsetAsync(base + q.offsets, 0, 4, p->stream))) return e;
        if ((e = d2d(q.offsets + 4, b.offsetsOff, (size_t)b.n * 4))) return e;
        if ((e = d2d(q.values, b.valuesOff, (size_t)b.dataLen))) return e;
        if (b.hasNulls) {
            col.nulls = (const uint8_t*)(base + q.nulls);
            p->stats.kernel_launches += 1;
            return unpackNullBits(p->stream, (const uint8_t*)(base + b.bitsOff), b.n, (uint8_t*)(base + q.nulls));
        }
        return e;
    };
    std::vector<pgpu_column> cols((size_t)nblocks), dicts((size_t)nblocks);
    for (int c = 0; c < nblocks; c++) {
        const WireBlock& b = blocks[c];
        const Type& t = plan.input[c];
        if (!b.child) {
            CU(decode(b, at[c], t, cols[c]));
            continue;
        }
        CU(decode(*b.child, atChild[c], t, dicts[c]));
        pgpu_column& col = cols[c];
        memset(&col, 0, sizeof col);
        col.encoding = b.kind == WireBlock::RLE ? PGPU_ENC_RLE : PGPU_ENC_DICTIONARY;
        col.type_kind = (int32_t)t.kind;
        col.precision = t.precision;
        col.scale = t.scale;
        col.position_count = b.n;
        col.dictionary = &dicts[c];
        if (b.kind == WireBlock::DICTIONARY) {
            CU(d2d(at[c].values, b.valuesOff, (size_t)b.n * 4));
            col.ids = (const int32_t*)(base + at[c].values);
        }
    }
    st = addPageCommon(p, cols.data(), nblocks, position_count, true, 0);
    return st ? st : updatePartialFill(p);
}

// host threads per chunk for the run-header walk: PGPU_PARQUET_THREADS, else the cores shared by the chunks walked at the same time
static int parquetHostThreads(int concurrentChunks) {
    if (const char* e = getenv("PGPU_PARQUET_THREADS")) return std::max(1, atoi(e));
    const unsigned hw = std::thread::hardware_concurrency();
    return (int)std::max(1u, std::min(16u, (hw ? hw : 1u) / (unsigned)std::max(concurrentChunks, 1)));
}

// ---- Parquet column chunks ------------------------------------------------------------------------------------------------
// Stands in for ParquetReader.nextPage / readBlock -> PrimitiveColumnReader.readPrimitive
// (lib/trino-parquet/src/main/java/io/trino/parquet/reader/ParquetReader.java:211-268, PrimitiveColumnReader.java:148-175):
// one row group's column chunks become the Blocks of one input Page. The host reads page and run headers only
// (parquet_host.h); the chunk bytes go to the device in one DMA per column and are expanded there (parquet.cu, serde.cu).
// A chunk whose data pages are all dictionary encoded becomes a DictionaryBlock (ids + the dictionary page's values, a NULL
// entry appended when the column has NULLs) — what the staged Q1 kernel consumes without materialising the values.
pgpu_status pgpu_add_parquet_row_group(pgpu_pipeline* p, const pgpu_parquet_chunk* chunks, int32_t nchunks, int64_t rows) {
    if (!p || (!chunks && nchunks)) return PGPU_INVALID_ARGUMENT;
    if (p->device < 0) return fail(p, PGPU_INVALID_STATE, "pipeline was created compile-only (device < 0)");
    if (p->finishing) return fail(p, PGPU_INVALID_STATE, "Operator is already finishing");
    if (!pgpu_needs_input(p)) return fail(p, PGPU_INVALID_STATE, "Operator does not need input");
    CU(cudaSetDevice(p->device));
    PqDecoded dec;
    std::string why;
    pgpu_status st = pqDecodeRowGroup(*p->plan, chunks, nchunks, rows, p->stream, p->dErr, parquetHostThreads(nchunks),
                                      [&](size_t need) -> char* { return ensureDev(p, p->parquetArena, need) == PGPU_OK ? (char*)p->parquetArena.ptr : nullptr; }, dec, why);
    p->stats.h2d_bytes += dec.h2dBytes;
    p->stats.kernel_launches += dec.launches;
    if (st) return why.empty() ? st : fail(p, st, why);
    if (rows == 0) return PGPU_OK;  // an empty row group adds nothing
    st = addPageCommon(p, dec.cols.data(), nchunks, rows, true, 0);
    return st ? st : updatePartialFill(p);
}

// Host-only view of what pgpu_add_parquet_row_group reads from a chunk (no GPU needed): info[0] data pages, [1] dictionary
// entries (-1: no dictionary page), [2] 1 if every data page is dictionary encoded, [3] non-null values, [4] definition level
// runs kept (0 when the chunk has no NULL), [5] value runs (ids / booleans), [6] 1 if some page is dictionary encoded, [7] rows.
// page headers + run headers of one chunk, after its pages have been decompressed if the chunk says they are compressed
static bool parseChunkOf(const pgpu_parquet_chunk* chunk, pq::Chunk& k, std::string& why) {
    std::vector<uint8_t> inflated;
    const uint8_t* d = chunk->data;
    size_t len = chunk->len;
    if (chunk->codec != pq::UNCOMPRESSED) {
        if (!pq::inflateChunk(chunk->data, chunk->len, chunk->codec, chunk->num_values, inflated, why)) return false;
        d = inflated.data();
        len = inflated.size();
    }
    return pq::parseChunk(d, len, chunk->physical_type, chunk->max_definition_level, chunk->num_values, k, why, chunk->type_length, parquetHostThreads(1));
}

pgpu_status pgpu_parquet_inspect_chunk(const pgpu_parquet_chunk* chunk, int64_t* info, char* err, size_t err_len) {
    if (!chunk || !info) return PGPU_INVALID_ARGUMENT;
    pq::Chunk k;
    std::string why;
    if (!parseChunkOf(chunk, k, why)) {
        if (err && err_len) {
            strncpy(err, why.c_str(), err_len - 1);
            err[err_len - 1] = 0;
        }
        return why.rfind("not supported", 0) == 0 ? PGPU_NOT_SUPPORTED : PGPU_INVALID_ARGUMENT;
    }
    info[0] = (int64_t)k.pages.size();
    info[1] = k.hasDictionary ? k.dictionaryCount : -1;
    info[2] = k.allDictionary ? 1 : 0;
    info[3] = k.nonNull;
    info[4] = (int64_t)k.levelRuns.size();
    info[5] = (int64_t)k.valueRuns.size();
    info[6] = k.anyDictionary ? 1 : 0;
    info[7] = k.rows;
    return PGPU_OK;
}

// The run tables pgpu_add_parquet_row_group hands to the device kernels (no GPU needed; used by the CPU test-suite to check them
// against the oracle): which = 0 definition levels, 1 values (dictionary ids / booleans), 2 data pages as
// {start = firstValue, src = byte offset of the values, count = non-null values, widthPacked = dictionary encoded ? 1 : 0}.
// Each entry = {int64 start, uint64 src, uint32 count, uint32 width | packed << 31}. *n = entries available.
pgpu_status pgpu_parquet_chunk_runs(const pgpu_parquet_chunk* chunk, int32_t which, void* out, int64_t cap, int64_t* n) {
    if (!chunk || !n) return PGPU_INVALID_ARGUMENT;
    pq::Chunk k;
    std::string why;
    if (!parseChunkOf(chunk, k, why)) return why.rfind("not supported", 0) == 0 ? PGPU_NOT_SUPPORTED : PGPU_INVALID_ARGUMENT;
    std::vector<pq::Run> pages;
    for (const pq::DataPage& pg : k.pages) pages.push_back(pq::Run{pg.firstValue, (uint64_t)pg.valuesOff, (uint32_t)pg.nonNull, pg.dictionaryEncoded ? 1u : 0u});
    const std::vector<pq::Run>& v = which == 0 ? k.levelRuns : which == 1 ? k.valueRuns : pages;
    *n = (int64_t)v.size();
    if (out && cap > 0) memcpy(out, v.data(), (size_t)std::min<int64_t>(cap, *n) * sizeof(pq::Run));
    return PGPU_OK;
}

pgpu_status pgpu_finish(pgpu_pipeline* p) {
    if (!p) return PGPU_INVALID_ARGUMENT;
    p->finishing = true;
    return p->device >= 0 ? flushPending(p) : PGPU_OK;  // the last (partly filled) super-page starts now, not at getOutput
}

int32_t pgpu_is_finished(const pgpu_pipeline* p) { return p && p->finished ? 1 : 0; }

enum class OutputTo { HOST, DEVICE, WIRE };

// COMPACT: the kernel of one input page has finished and passed `k` rows, written behind the mergeRows buffered ones.
// MergePages.java:112-151: a page with at least merge-min-rows rows or merge-min-bytes bytes goes out as it is, after the
// buffer if that holds rows (two output pages); a smaller one joins the buffer, which goes out once it holds merge-max-bytes.
static void compactPageDone(pgpu_pipeline* p, int64_t k) {
    const Plan& plan = *p->plan;
    if (!p->outRowBytes)
        for (auto& d : p->layout.out) p->outRowBytes += std::max(d.stride, 1) + 1;  // values + the null flag, as Block.getSizeInBytes counts them
    const bool merging = plan.mergeMinRows > 0 || plan.mergeMinBytes > 0;
    const bool big = !merging || k >= plan.mergeMinRows || k * p->outRowBytes >= plan.mergeMinBytes;
    if (big) {
        p->outBoundary = p->mergeRows;
        p->outRowsTotal = p->mergeRows + k;
        p->outRowsServed = 0;
        p->mergeRows = 0;
        p->outputReady = true;  // an empty page too: the caller sees a page of 0 rows, as before
        return;
    }
    p->mergeRows += k;
    p->stats.pages_merged++;
    if (p->mergeRows * p->outRowBytes >= plan.mergeMaxBytes) {  // PageBuilder.isFull
        p->outRowsTotal = p->mergeRows;
        p->outRowsServed = p->outBoundary = p->mergeRows = 0;
        p->outputReady = true;
        return;
    }
    p->outputReady = false;  // needs more data
}

static pgpu_status serializeOutput(pgpu_pipeline* p, int64_t r0, int64_t n);

static pgpu_status getOutputImpl(pgpu_pipeline* p, pgpu_column* out_cols, int32_t max_cols, int32_t* ncols, int64_t* rows, OutputTo to) {
    const bool toHost = to == OutputTo::HOST;
    if (!p || !ncols || !rows) return PGPU_INVALID_ARGUMENT;
    *rows = -1;
    *ncols = 0;
    if (p->finished) return PGPU_OK;
    if (p->device < 0) return fail(p, PGPU_INVALID_STATE, "pipeline was created compile-only (device < 0)");
    const GenResult& g = p->layout;
    const Plan& plan = *p->plan;
    CU(cudaSetDevice(p->device));
    if (g.mode == KernelMode::COMPACT) {
        if (p->compactPending) {
            CU(cudaStreamSynchronize(p->stream));
            pgpu_status st = checkDeviceErrors(p);
            if (st) return st;
            p->compactPending = false;
            compactPageDone(p, (int64_t)p->hostScalars[0]);
        }
        if (!p->outputReady) {
            if (!p->finishing) return PGPU_OK;
            if (p->mergeRows == 0) {
                p->finished = true;
                return PGPU_OK;
            }
            // MergePages.java:125-131: input finished, flush what the builder holds
            p->outRowsTotal = p->mergeRows;
            p->outRowsServed = p->outBoundary = p->mergeRows = 0;
            p->outputReady = true;
        }
    } else {
        if (!p->outputReady) {
            const bool flush = isPartialFlushDue(p);
            if (!p->finishing && !flush) return PGPU_OK;  // only flush if we are finishing or the builder is full
            const bool global = plan.groupChannels.empty();
            if (!global && !p->sawInput) {  // HashAggregationOperator.getOutput: no input, no default output for grouped aggregation
                p->finished = true;
                return PGPU_OK;
            }
            pgpu_status st = buildAggOutput(p);
            if (st) return st;
        }
    }
    if (to != OutputTo::WIRE && (max_cols < g.NO || !out_cols)) return fail(p, PGPU_INVALID_ARGUMENT, "out_cols too small: need " + std::to_string(g.NO));
    int64_t remaining = p->outRowsTotal - p->outRowsServed;
    int64_t n = remaining;
    if (plan.maxOutputRows > 0) n = std::min<int64_t>(n, plan.maxOutputRows);
    const int64_t r0 = p->outRowsServed;
    if (g.mode == KernelMode::COMPACT && r0 < p->outBoundary) n = std::min<int64_t>(n, p->outBoundary - r0);
    if (remaining <= 0 && g.mode != KernelMode::COMPACT) {
        // nothing (left) to serve
        p->outputReady = false;
        if (p->finishing) p->finished = true;
        else resetAggState(p);
        return PGPU_OK;
    }
    // small results: the whole output arena in one copy (a copy call costs more than these bytes take on the wire)
    const bool oneCopy = toHost && n > 0 && p->outLayoutBytes <= (256u << 10);
    if (oneCopy) {
        pgpu_status st = ensureHost(p, p->outHostArena, p->outLayoutBytes);
        if (st) return st;
        CU(cudaMemcpyAsync(p->outHostArena.ptr, p->outArena.ptr, p->outLayoutBytes, cudaMemcpyDeviceToHost, p->stream));
        p->stats.d2h_bytes += (int64_t)p->outLayoutBytes;
    }
    if (to == OutputTo::WIRE) {
        pgpu_status st = serializeOutput(p, r0, n);
        if (st) return st;
    }
    for (int c = 0; c < g.NO && to != OutputTo::WIRE; c++) {
        const OutColDesc& d = g.out[c];
        pgpu_column& oc = out_cols[c];
        memset(&oc, 0, sizeof oc);
        oc.encoding = PGPU_ENC_PLAIN;
        oc.type_kind = (int32_t)d.type.kind;
        oc.precision = d.type.precision;
        oc.scale = d.type.scale;
        oc.position_count = n;
        const char* dv = (const char*)p->outDev[c].ptr + (size_t)r0 * d.stride;
        const uint8_t* dn = (const uint8_t*)p->outDevNulls[c].ptr + r0;
        if (!toHost) {
            if (d.varWidth && !(d.lenPrefixed && d.stride == 16)) return fail(p, PGPU_NOT_SUPPORTED, "not supported on the GPU path: device-resident var-width output");
            if (d.varWidth) oc.encoding = PGPU_ENC_VARCELL;
            oc.values = dv;
            oc.nulls = dn;
            continue;
        }
        if (oneCopy) {
            oc.values = (char*)p->outHostArena.ptr + p->outOffV[c] + (size_t)r0 * d.stride;
            oc.nulls = (const uint8_t*)p->outHostArena.ptr + p->outOffN[c] + r0;
            continue;
        }
        pgpu_status st = ensureHost(p, p->outHost[c], (size_t)std::max<int64_t>(n, 1) * std::max(d.stride, 1));
        if (st) return st;
        st = ensureHost(p, p->outHostNulls[c], (size_t)std::max<int64_t>(n, 1));
        if (st) return st;
        if (n > 0) {
            CU(cudaMemcpyAsync(p->outHost[c].ptr, dv, (size_t)n * d.stride, cudaMemcpyDeviceToHost, p->stream));
            CU(cudaMemcpyAsync(p->outHostNulls[c].ptr, dn, (size_t)n, cudaMemcpyDeviceToHost, p->stream));
            p->stats.d2h_bytes += n * d.stride + n;
        }
        oc.values = p->outHost[c].ptr;
        oc.nulls = (const uint8_t*)p->outHostNulls[c].ptr;
    }
    if (toHost) {
        CU(cudaStreamSynchronize(p->stream));
        // var-width cells -> VariableWidthBlock layout (offsets + bytes), in place
        for (int c = 0; c < g.NO; c++) {
            const OutColDesc& d = g.out[c];
            if (!d.varWidth) continue;
            pgpu_status st = ensureHost(p, p->outHostAux[c], (size_t)(n + 1) * 4);
            if (st) return st;
            int32_t* offs = (int32_t*)p->outHostAux[c].ptr;
            unsigned char* cells = (unsigned char*)const_cast<void*>(out_cols[c].values);
            const uint8_t* nulls = out_cols[c].nulls;
            int32_t w = 0;
            offs[0] = 0;
            for (int64_t i = 0; i < n; i++) {
                const unsigned char* cell = cells + (size_t)i * d.stride;
                int len = d.lenPrefixed ? cell[0] : d.stride;
                if (nulls[i]) len = 0;
                memmove(cells + w, cell + (d.lenPrefixed ? 1 : 0), (size_t)len);  // w <= i*stride: never overtakes the read cursor
                w += len;
                offs[i + 1] = w;
            }
            out_cols[c].offsets = offs;
        }
    }
    *ncols = g.NO;
    *rows = n;
    p->outRowsServed += n;
    if (g.mode == KernelMode::COMPACT) {
        // a page that passes more rows than max-output-rows goes out in several calls; needsInput stays false in between
        if (p->outRowsServed >= p->outRowsTotal) p->outputReady = false;
    } else if (p->outRowsServed >= p->outRowsTotal) {
        p->outputReady = false;
        if (p->finishing) p->finished = true;
        else resetAggState(p);  // partial flush: start a fresh table (HashAggregationOperator.closeAggregationBuilder)
    }
    return PGPU_OK;
}

pgpu_status pgpu_get_output(pgpu_pipeline* p, pgpu_column* out_cols, int32_t max_cols, int32_t* ncols, int64_t* rows) {
    return getOutputImpl(p, out_cols, max_cols, ncols, rows, OutputTo::HOST);
}

pgpu_status pgpu_get_output_device(pgpu_pipeline* p, pgpu_column* out_cols, int32_t max_cols, int32_t* ncols, int64_t* rows) {
    return getOutputImpl(p, out_cols, max_cols, ncols, rows, OutputTo::DEVICE);
}

// Rows [r0, r0 + n) of the output buffers -> the payload PagesSerdeUtil.writeRawPage would produce for the same blocks, in
// p->serdeHost. The device does the packing (null bits, non-null-only values, end offsets + bytes of var-width cells) into
// aligned scratch; every piece is then copied straight to its (unaligned) place in the pinned payload, next to the block
// headers the host writes. Two synchronisations: one to learn the non-null counts / byte totals the layout depends on, one
// at the end.
static pgpu_status serializeOutput(pgpu_pipeline* p, int64_t r0, int64_t n) {
    const GenResult& g = p->layout;
    const int NO = g.NO;
    struct Col {
        const char* values;
        const uint8_t* nulls;
        size_t bits = 0, counts = 0, compact = 0, offsets = 0, byteCounts = 0;  // scratch offsets
        uint32_t total = 0;                                     // non-null positions / data bytes
    };
    std::vector<Col> cols((size_t)NO);
    const size_t nblocks = packScratchWords(n);
    size_t need = 0;
    auto reserve = [&](size_t bytes) {
        size_t at = need;
        need += align256(std::max<size_t>(bytes, 16));
        return at;
    };
    for (int c = 0; c < NO; c++) {
        const OutColDesc& d = g.out[c];
        Col& k = cols[c];
        k.values = (const char*)p->outDev[c].ptr + (size_t)r0 * d.stride;
        k.nulls = (const uint8_t*)p->outDevNulls[c].ptr + r0;
        k.bits = reserve((size_t)(n + 7) / 8);
        k.counts = reserve(nblocks * 4);
        k.compact = reserve((size_t)n * std::max(d.stride, 1));
        if (d.varWidth) {
            k.offsets = reserve((size_t)n * 4);
            k.byteCounts = reserve(nblocks * 4);
        } else if (d.stride != 1 && d.stride != 4 && d.stride != 8 && d.stride != 16)
            return fail(p, PGPU_NOT_SUPPORTED, "not supported on the GPU path: wire format of a " + std::to_string(d.stride) + "-byte fixed-width channel");
    }
    pgpu_status st = ensureDev(p, p->serdeOutArena, need);
    if (st) return st;
    st = ensureHost(p, p->serdeTotals, (size_t)NO * 2 * 4);
    if (st) return st;
    char* base = (char*)p->serdeOutArena.ptr;
    uint32_t* totals = (uint32_t*)p->serdeTotals.ptr;  // [c]: non-null count, [NO + c]: var-width bytes
    for (int c = 0; c < NO && n > 0; c++) {
        const OutColDesc& d = g.out[c];
        const Col& k = cols[c];
        uint32_t* counts = (uint32_t*)(base + k.counts);
        CU(packNullBits(p->stream, k.nulls, n, (uint8_t*)(base + k.bits), counts));
        CU(cudaMemcpyAsync(&totals[c], counts + (nblocks - 1), 4, cudaMemcpyDeviceToHost, p->stream));
        p->stats.kernel_launches += 2;
        if (!d.varWidth) continue;
        counts = (uint32_t*)(base + k.byteCounts);
        CU(cellByteCounts(p->stream, (const uint8_t*)k.values, k.nulls, n, d.stride, d.lenPrefixed, counts));
        CU(cudaMemcpyAsync(&totals[NO + c], counts + (nblocks - 1), 4, cudaMemcpyDeviceToHost, p->stream));
        p->stats.kernel_launches += 2;
    }
    CU(cudaStreamSynchronize(p->stream));
    // layout of the payload
    auto encodingName = [](const OutColDesc& d) -> const char* {
        if (d.varWidth) return "VARIABLE_WIDTH";
        return d.stride == 8 ? "LONG_ARRAY" : d.stride == 4 ? "INT_ARRAY" : d.stride == 1 ? "BYTE_ARRAY" : "INT128_ARRAY";
    };
    size_t len = 4;
    for (int c = 0; c < NO; c++) {
        const OutColDesc& d = g.out[c];
        const uint32_t nonNull = n > 0 ? totals[c] : 0;
        const bool hasNulls = nonNull != (uint32_t)n;
        len += 4 + strlen(encodingName(d)) + 4;
        if (d.varWidth) len += (size_t)n * 4 + 1 + (hasNulls ? (size_t)(n + 7) / 8 : 0) + 4 + (n > 0 ? totals[NO + c] : 0);
        else len += 1 + (hasNulls ? (size_t)(n + 7) / 8 + 4 : 0) + (size_t)nonNull * d.stride;
    }
    st = ensureHost(p, p->serdeHost, len);
    if (st) return st;
    uint8_t* out = (uint8_t*)p->serdeHost.ptr;
    size_t pos = 0;
    auto putInt = [&](int32_t v) {
        memcpy(out + pos, &v, 4);
        pos += 4;
    };
    auto d2h = [&](const void* src, size_t bytes) -> cudaError_t {
        cudaError_t e = bytes ? cudaMemcpyAsync(out + pos, src, bytes, cudaMemcpyDeviceToHost, p->stream) : cudaSuccess;
        pos += bytes;
        p->stats.d2h_bytes += (int64_t)bytes;
        return e;
    };
    putInt(NO);  // PagesSerdeUtil.writeRawPage: channel count, then the blocks
    for (int c = 0; c < NO; c++) {
        const OutColDesc& d = g.out[c];
        const Col& k = cols[c];
        const uint32_t nonNull = n > 0 ? totals[c] : 0;
        const bool hasNulls = nonNull != (uint32_t)n;
        const char* name = encodingName(d);
        putInt((int32_t)strlen(name));
        memcpy(out + pos, name, strlen(name));
        pos += strlen(name);
        putInt((int32_t)n);
        if (d.varWidth) {  // spi/block/VariableWidthBlockEncoding.java:37-58: end offsets, nulls, total length, bytes
            const uint32_t bytes = n > 0 ? totals[NO + c] : 0;
            if (n > 0) {
                CU(packCells(p->stream, (const uint8_t*)k.values, k.nulls, n, d.stride, d.lenPrefixed, (const uint32_t*)(base + k.byteCounts),
                             (int32_t*)(base + k.offsets), (uint8_t*)(base + k.compact)));
                p->stats.kernel_launches += 1;
            }
            CU(d2h(base + k.offsets, (size_t)n * 4));
            out[pos++] = hasNulls ? 1 : 0;
            if (hasNulls) CU(d2h(base + k.bits, (size_t)(n + 7) / 8));
            putInt((int32_t)bytes);
            CU(d2h(base + k.compact, bytes));
        } else {  // spi/block/LongArrayBlockEncoding.java:39-60 and siblings: nulls, [non-null count], non-null values
            out[pos++] = hasNulls ? 1 : 0;
            if (!hasNulls) {
                CU(d2h(k.values, (size_t)n * d.stride));
                continue;
            }
            CU(d2h(base + k.bits, (size_t)(n + 7) / 8));
            putInt((int32_t)nonNull);
            CU(compactNonNull(p->stream, k.nulls, n, k.values, base + k.compact, d.stride, (const uint32_t*)(base + k.counts)));
            p->stats.kernel_launches += 1;
            CU(d2h(base + k.compact, (size_t)nonNull * d.stride));
        }
    }
    if (pos != len) return fail(p, PGPU_CUDA_ERROR, "internal: wire layout mismatch");
    CU(cudaStreamSynchronize(p->stream));
    p->serdeLen = len;
    return PGPU_OK;
}

pgpu_status pgpu_get_serialized_output(pgpu_pipeline* p, const uint8_t** data, size_t* len, int64_t* rows) {
    if (!p || !data || !len || !rows) return PGPU_INVALID_ARGUMENT;
    *data = nullptr;
    *len = 0;
    int32_t nc = 0;
    pgpu_status st = getOutputImpl(p, nullptr, 0, &nc, rows, OutputTo::WIRE);
    if (st || *rows < 0) return st;
    *data = (const uint8_t*)p->serdeHost.ptr;
    *len = p->serdeLen;
    return PGPU_OK;
}

pgpu_status pgpu_sync(pgpu_pipeline* p) {
    if (!p) return PGPU_INVALID_ARGUMENT;
    if (p->device < 0) return PGPU_OK;
    CU(cudaSetDevice(p->device));
    {
        pgpu_status st = flushPending(p);
        if (st) return st;
    }
    CU(cudaStreamSynchronize(p->stream));
    for (auto& s : p->slots) s.inFlight = false;
    return checkDeviceErrors(p);
}

pgpu_status pgpu_reset(pgpu_pipeline* p) {
    if (!p) return PGPU_INVALID_ARGUMENT;
    p->pend.rows = 0;  // rows waiting in the staging slot are dropped with the rest of the state
    p->extraOutRows = 0;
    if (p->device >= 0) {
        CU(cudaSetDevice(p->device));
        resetAggState(p);
        CU(cudaStreamSynchronize(p->stream));
        for (auto& s : p->slots) s.inFlight = false;
    }
    p->runRows = 0;  // (the streaming decision stays: STREAM_OFF once the keys were seen out of order)
    p->finishing = p->finished = p->sawInput = p->outputReady = p->compactPending = false;
    p->outRowsTotal = p->outRowsServed = p->mergeRows = p->outBoundary = 0;
    int64_t cap = p->stats.table_capacity, bucketed = p->stats.table_bucketed;
    p->stats = pgpu_stats{};
    p->stats.table_capacity = cap;
    p->stats.table_bucketed = bucketed;
    p->stats.bucket_offset = (int64_t)p->bucketOff;
    p->err.clear();
    return PGPU_OK;
}

int32_t pgpu_output_channel_count(const pgpu_pipeline* p) { return p ? (int32_t)p->plan->outputTypes().size() : 0; }

pgpu_status pgpu_output_channel_type(const pgpu_pipeline* p, int32_t channel, int32_t* type_kind, int32_t* precision, int32_t* scale) {
    if (!p) return PGPU_INVALID_ARGUMENT;
    std::vector<Type> t = p->plan->outputTypes();
    if (channel < 0 || channel >= (int)t.size()) return PGPU_INVALID_ARGUMENT;
    if (type_kind) *type_kind = (int32_t)t[channel].kind;
    if (precision) *precision = t[channel].precision;
    if (scale) *scale = t[channel].scale;
    return PGPU_OK;
}

const char* pgpu_generated_source(const pgpu_pipeline* p) { return p ? p->lastSource.c_str() : ""; }

pgpu_status pgpu_get_stats(const pgpu_pipeline* p, pgpu_stats* out) {
    if (!p || !out) return PGPU_INVALID_ARGUMENT;
    *out = p->stats;
    return PGPU_OK;
}

pgpu_status pgpu_set_timing(pgpu_pipeline* p, int32_t enabled) {
    if (!p) return PGPU_INVALID_ARGUMENT;
    p->timing = enabled != 0;
    return PGPU_OK;
}

void* pgpu_stream(const pgpu_pipeline* p) { return p ? (void*)p->stream : nullptr; }

pgpu_status pgpu_compile_signature(pgpu_pipeline* p, const int32_t* encodings, const uint8_t* nullable, int32_t ncols) {
    if (!p || !encodings || !nullable) return PGPU_INVALID_ARGUMENT;
    if (ncols != (int)p->plan->input.size()) return fail(p, PGPU_INVALID_ARGUMENT, "signature length differs from the plan's input channels");
    std::vector<ColSig> sig(ncols);
    for (int c = 0; c < ncols; c++) {
        sig[c].encoding = encodings[c];
        sig[c].nullable = nullable[c] != 0;
    }
    Variant* v = nullptr;
    return getVariant(p, sig, &v);
}

void* pgpu_host_alloc(size_t bytes) {
    void* q = nullptr;
    if (cudaMallocHost(&q, bytes ? bytes : 1) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return q;
}

void pgpu_host_free(void* ptr) {
    if (ptr) cudaFreeHost(ptr);
}

void* pgpu_device_alloc(int32_t device, size_t bytes) {
    void* q = nullptr;
    if (cudaSetDevice(device) != cudaSuccess || cudaMalloc(&q, bytes ? bytes : 1) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return q;
}

void pgpu_device_free(int32_t device, void* ptr) {
    if (!ptr) return;
    cudaSetDevice(device);
    cudaFree(ptr);
}

pgpu_status pgpu_partition_output_device(pgpu_pipeline* p, int32_t nparts, pgpu_column* out_cols, int32_t max_cols, int32_t* ncols,
                                         int64_t* part_row_offsets, int64_t* part_row_counts) {
    if (!p || !ncols || !part_row_offsets || !part_row_counts) return PGPU_INVALID_ARGUMENT;
    if (p->device < 0) return fail(p, PGPU_INVALID_STATE, "pipeline was created compile-only (device < 0)");
    const Plan& plan = *p->plan;
    const GenResult& g = p->layout;
    if (!plan.hasAgg || plan.groupChannels.empty()) return fail(p, PGPU_NOT_SUPPORTED, "not supported on the GPU path: partitioned output of a pipeline without grouping keys");
    if (nparts < 1 || nparts > kMaxParts) return fail(p, PGPU_INVALID_ARGUMENT, "partition count must be in [1, 64]");
    if (!p->finishing) return fail(p, PGPU_INVALID_STATE, "partitioned output needs finish() first");
    const int nkeys = (int)plan.groupChannels.size();
    if (nkeys > kMaxPartKeys) return fail(p, PGPU_NOT_SUPPORTED, "not supported on the GPU path: more than 8 partitioning channels");
    if (max_cols < g.NO || !out_cols) return fail(p, PGPU_INVALID_ARGUMENT, "out_cols too small: need " + std::to_string(g.NO));
    PartKeys keys{};
    keys.nkeys = nkeys;
    keys.nparts = nparts;
    for (int k = 0; k < nkeys; k++) {
        const Type& t = g.out[k].type;
        if (t.kind == TypeKind::BIGINT) keys.k[k].kind = PART_KEY_I64;
        else if (t.kind == TypeKind::INTEGER || t.kind == TypeKind::DATE) keys.k[k].kind = PART_KEY_I32;
        else if (t.kind == TypeKind::DOUBLE) keys.k[k].kind = PART_KEY_F64;
        else return fail(p, PGPU_NOT_SUPPORTED, "not supported on the GPU path: hash partitioning on " + t.str() + " keys");
    }
    for (int c = 0; c < g.NO; c++)
        if (g.out[c].varWidth && g.out[c].lenPrefixed) return fail(p, PGPU_NOT_SUPPORTED, "not supported on the GPU path: device-resident var-width output");
    CU(cudaSetDevice(p->device));
    for (int i = 0; i <= nparts; i++) part_row_offsets[i] = 0;
    for (int i = 0; i < nparts; i++) part_row_counts[i] = 0;
    *ncols = g.NO;
    int64_t rows = 0;
    if (!p->finished && (p->outputReady || p->sawInput)) {
        if (!p->outputReady) {
            pgpu_status st = buildAggOutput(p);
            if (st) return st;
        }
        rows = p->outRowsTotal - p->outRowsServed;
    }
    const int64_t r0 = p->outRowsServed;
    if (!p->hostPart) CU(cudaMallocHost((void**)&p->hostPart, 2 * kMaxParts * sizeof(unsigned long long)));
    const int64_t rr = std::max<int64_t>(rows, 1);
    pgpu_status st;
    if ((st = ensureDev(p, p->partPid, (size_t)rr))) return st;
    if ((st = ensureDev(p, p->partDst, (size_t)rr * 8))) return st;
    if ((st = ensureDev(p, p->partCounts, 2 * kMaxParts * sizeof(unsigned long long)))) return st;
    // every partition starts at a multiple of 16 rows: each partition of each column is a 16-byte aligned device block
    // that a FINAL pipeline (local or a peer's, over NVLink) accepts as it is. Sized by the output capacity so that the
    // allocation (and a peer's mapping of it) is stable from step to step.
    const int64_t capRows = (std::max(rr, p->outCapacityRows) + 15) / 16 * 16 + 16 * (int64_t)kMaxParts;
    size_t need = 0;
    std::vector<size_t> offV(g.NO), offN(g.NO);
    for (int c = 0; c < g.NO; c++) {
        offV[c] = need;
        need += align256((size_t)capRows * std::max(g.out[c].stride, 1));
        offN[c] = need;
        need += align256((size_t)capRows);
    }
    if (p->partArena.cap < need) {
        CU(cudaStreamSynchronize(p->stream));
        if (p->partArena.ptr) p->partRetired.push_back(p->partArena.ptr);
        p->partArena = DeviceBuf{};
        const size_t want = (need + (2u << 20) - 1) / (2u << 20) * (2u << 20);
        CU(cudaMalloc(&p->partArena.ptr, want));
        p->partArena.cap = want;
    }
    p->partDev.assign(g.NO, nullptr);
    p->partDevNulls.assign(g.NO, nullptr);
    for (int c = 0; c < g.NO; c++) {
        p->partDev[c] = (char*)p->partArena.ptr + offV[c];
        p->partDevNulls[c] = (char*)p->partArena.ptr + offN[c];
    }
    if (rows > 0) {
        for (int k = 0; k < nkeys; k++) {
            keys.k[k].values = (const char*)p->outDev[k].ptr + (size_t)r0 * g.out[k].stride;
            keys.k[k].nulls = (const uint8_t*)p->outDevNulls[k].ptr + r0;
        }
        unsigned long long* dCounts = (unsigned long long*)p->partCounts.ptr;
        unsigned long long* dCursors = dCounts + kMaxParts;
        CU(cudaMemsetAsync(dCounts, 0, 2 * kMaxParts * sizeof(unsigned long long), p->stream));
        CU(partitionCount(p->stream, p->numSMs, keys, rows, (uint8_t*)p->partPid.ptr, dCounts));
        CU(cudaMemcpyAsync(p->hostPart, dCounts, kMaxParts * sizeof(unsigned long long), cudaMemcpyDeviceToHost, p->stream));
        CU(cudaStreamSynchronize(p->stream));
        unsigned long long acc = 0, real = 0;
        for (int i = 0; i < nparts; i++) {
            p->hostPart[kMaxParts + i] = acc;
            part_row_offsets[i] = (int64_t)acc;
            part_row_counts[i] = (int64_t)p->hostPart[i];
            real += p->hostPart[i];
            acc = (acc + p->hostPart[i] + 15) / 16 * 16;
        }
        part_row_offsets[nparts] = (int64_t)acc;
        if ((int64_t)real != rows) return fail(p, PGPU_CUDA_ERROR, "internal: partition counts do not add up");
        CU(cudaMemcpyAsync(dCursors, p->hostPart + kMaxParts, kMaxParts * sizeof(unsigned long long), cudaMemcpyHostToDevice, p->stream));
        CU(partitionAssign(p->stream, p->numSMs, nparts, rows, (const uint8_t*)p->partPid.ptr, dCursors, (int64_t*)p->partDst.ptr));
        for (int c = 0; c < g.NO; c++) {
            const int stride = g.out[c].stride;
            CU(partitionScatter(p->stream, p->numSMs, rows, (const int64_t*)p->partDst.ptr, (const char*)p->outDev[c].ptr + (size_t)r0 * stride, p->partDev[c], stride));
            CU(partitionScatter(p->stream, p->numSMs, rows, (const int64_t*)p->partDst.ptr, (const uint8_t*)p->outDevNulls[c].ptr + r0, p->partDevNulls[c], 1));
        }
        p->stats.kernel_launches += 2 + 2 * g.NO;
        CU(cudaStreamSynchronize(p->stream));
    }
    for (int c = 0; c < g.NO; c++) {
        pgpu_column& oc = out_cols[c];
        memset(&oc, 0, sizeof oc);
        oc.encoding = PGPU_ENC_PLAIN;
        oc.type_kind = (int32_t)g.out[c].type.kind;
        oc.precision = g.out[c].type.precision;
        oc.scale = g.out[c].type.scale;
        oc.position_count = part_row_offsets[nparts];
        oc.values = p->partDev[c];
        oc.nulls = (const uint8_t*)p->partDevNulls[c];
    }
    p->outRowsServed = p->outRowsTotal;
    p->outputReady = false;
    p->finished = true;
    return PGPU_OK;
}

// ---- peer access over NVLink / NVSwitch (one process per GPU) ------------------------------------------------------------
namespace {
typedef int (*AddressRangeFn)(unsigned long long* base, size_t* size, unsigned long long ptr);
AddressRangeFn addressRangeFn() {
    static AddressRangeFn fn = []() -> AddressRangeFn {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) return nullptr;
        return (AddressRangeFn)f;
    }();
    return fn;
}
}  // namespace

pgpu_status pgpu_ipc_export(int32_t device, const void* dev_ptr, uint8_t* handle, int64_t* offset) {
    if (!dev_ptr || !handle || !offset) return PGPU_INVALID_ARGUMENT;
    static_assert(sizeof(cudaIpcMemHandle_t) == PGPU_IPC_HANDLE_BYTES, "handle size");
    if (cudaSetDevice(device) != cudaSuccess) return PGPU_CUDA_ERROR;
    AddressRangeFn range = addressRangeFn();
    unsigned long long base = 0;
    size_t size = 0;
    if (!range || range(&base, &size, (unsigned long long)(uintptr_t)dev_ptr) != 0) return PGPU_CUDA_ERROR;
    cudaIpcMemHandle_t h;
    if (cudaIpcGetMemHandle(&h, (void*)(uintptr_t)base) != cudaSuccess) {
        cudaGetLastError();
        return PGPU_CUDA_ERROR;
    }
    memcpy(handle, &h, sizeof h);
    *offset = (int64_t)((unsigned long long)(uintptr_t)dev_ptr - base);
    return PGPU_OK;
}

pgpu_status pgpu_ipc_open(int32_t device, const uint8_t* handle, void** mapped_base) {
    if (!handle || !mapped_base) return PGPU_INVALID_ARGUMENT;
    if (cudaSetDevice(device) != cudaSuccess) return PGPU_CUDA_ERROR;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    cudaError_t e = cudaIpcOpenMemHandle(mapped_base, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        cudaGetLastError();
        pgpu::setLastGlobalError(std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e));
        return PGPU_CUDA_ERROR;
    }
    return PGPU_OK;
}

pgpu_status pgpu_ipc_close(int32_t device, void* mapped_base) {
    if (!mapped_base) return PGPU_INVALID_ARGUMENT;
    if (cudaSetDevice(device) != cudaSuccess || cudaIpcCloseMemHandle(mapped_base) != cudaSuccess) {
        cudaGetLastError();
        return PGPU_CUDA_ERROR;
    }
    return PGPU_OK;
}

int32_t pgpu_partition_of(const int64_t* key_bits, const uint8_t* is_null, const int32_t* kinds, int32_t nkeys, int32_t nparts) {
    if (!key_bits || !is_null || !kinds || nkeys < 0 || nparts < 1) return -1;
    return partitionOfHost(key_bits, is_null, kinds, nkeys, nparts);
}

}  // extern "C"

#include "comm.inc"
