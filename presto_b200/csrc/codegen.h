// codegen.h — Plan + page-encoding signature -> CUDA source for NVRTC.
//
// Replaces the reference's bytecode generators for this path: main/sql/gen/PageFunctionCompiler.java:219-365
// (projections), :459-544 (filter), main/sql/gen/RowExpressionCompiler.java:78-229 and the special-form generators,
// main/operator/aggregation/AccumulatorCompiler.java:261-330,490-569 (accumulator loops).
#pragma once
#include <string>
#include <vector>

#include "plan.h"

namespace pgpu {

// How one input channel arrives in a page. Kernel variants are specialised (and cached) per signature.
struct ColSig {
    int encoding = 0;        // pgpu_encoding
    bool nullable = false;   // PLAIN: valueIsNull present; DICTIONARY/RLE: the dictionary/value block has nulls
    bool operator==(const ColSig& o) const { return encoding == o.encoding && nullable == o.nullable; }
};

enum class KernelMode { COMPACT, AGG_GLOBAL, AGG_HASH };

struct OutColDesc {
    Type type;
    int stride = 0;          // bytes per row in the device output buffer
    bool varWidth = false;   // device buffer holds fixed-stride cells that the host turns into offsets + bytes
    bool lenPrefixed = false;// varchar key cells: byte 0 = length, bytes 1.. = data; else `stride` payload bytes
};

struct GenResult {
    std::string source;
    KernelMode mode = KernelMode::COMPACT;
    int NC = 0, KW = 0, NW = 0, SLOTW = 0, NO = 0;
    bool useDir = false;
    int dirCap = 0;
    int smemBytes = 0;         // dynamic shared memory of the page kernel
    int locBits = 0;           // table placement keeps 2^locBits neighbouring keys in consecutive slots (PGPU_LOCB)
    // staged variant (pgpu_k_page_staged): bulk-async copies of column tiles into a shared-memory ring
    bool staged = false;
    int stagedThreads = 544;   // 2 x 8 consumer warps + 1 producer warp
    int stagedTile = 1024;     // rows per stage
    int stageBytes = 0;        // bytes of one stage
    int stagedFixedSmem = 0;   // directory + warp tables + mbarriers in front of the stages
    std::vector<int> dcacheChannels;  // staged kernel: channels whose dictionary is decoded into shared memory
    int dcacheCap = 0;                // ... which needs dictionaries of at most this many entries
    double keyDomain = 0;      // upper bound of distinct group keys when the key types bound it (0 = unbounded)
    std::vector<int> cellKinds;  // NW entries (CK_*)
    std::vector<OutColDesc> out;
    std::vector<int> usedChannels;  // input channels the kernels dereference
    // bucketed table (8 neighbouring keys of one integer key field per slot; device/pgpu_device.cuh "bucketed slots")
    bool bucketsPossible = false;  // the plan qualifies: high cardinality, one integer key, run compaction fits
    bool buckets = false;          // this source was generated for the bucketed table: SLOTW = words per BUCKET
    bool rangeNullable = false;    // ... and that key may be NULL (the NULL group is always exchanged)
    bool rangeOk = false;          // one integer grouping key: the exchange may split groups by key range (pgpu_k_emit_range / pgpu_k_keyrange)
    int streamSub = 1;             // row quads per thread of pgpu_k_runs_count / _write (a warp unit = 32 x 4 streamSub rows)
    bool streamOk = false;         // streaming-runs kernels were generated (pgpu_k_runs_count / _write / _emit / _to_table)
    bool sweepOk = false;          // partition sweep kernels were generated (pgpu_k_part_count / _scatter / pgpu_k_page_sweep)
    int sweepRowBytes = 0;         // bytes the scatter moves per row (arrays of the used channels)
    int sweepStageWidth = 1;       // widest moved element: the scatter stages kPartTile elements of one array at a time
    int streamKeyChannel = -1;     // input channel the FIRST grouping field is read from as it is, an integer type (-1: none): the streaming-runs probe
    int bucketKeyChannel = -1;     // input channel the key is read from as it is (-1: the key is computed), for the clustering probe
};

struct GenOptions {
    bool buckets = false;  // generate the bucketed-table kernels (ignored unless GenResult::bucketsPossible)
};

// throws PlanError(PGPU_NOT_SUPPORTED) for expression forms that parse but have no GPU implementation
GenResult generate(const Plan& plan, const std::vector<ColSig>& sig, const GenOptions& opt = GenOptions());

}  // namespace pgpu
