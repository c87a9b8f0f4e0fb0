// exchange.h — device side of the partial -> final exchange: hash-partition the rows of a PARTIAL-step output page.
// Stands in for PartitionedOutputOperator's partition function (main/operator/PartitionedOutputOperator.java:56-133 ->
// HashGenerator.getPartition, main/operator/HashGenerator.java:24-35): partition = (rawHash & 0x7fff...) % partitionCount,
// rawHash = 31 * previous + hash(column) over the key channels (main/operator/InterpretedHashGenerator.java:66-74,
// main/operator/scalar/CombineHashFunction.java:26-29), NULL hashes to 0, initial value 0.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pgpu {

enum PartKeyKind { PART_KEY_I64 = 0, PART_KEY_I32 = 1, PART_KEY_F64 = 2 };

struct PartKey {
    const void* values;
    const uint8_t* nulls;  // 1 byte / row or nullptr
    int32_t kind;          // PartKeyKind
    int32_t pad;
};

constexpr int kMaxPartKeys = 8;
constexpr int kMaxParts = 64;

struct PartKeys {
    PartKey k[kMaxPartKeys];
    int32_t nkeys;
    int32_t nparts;
};

// pass 1: pid[row] = partition of the row; counts[p] += rows of partition p (counts zeroed by the caller)
cudaError_t partitionCount(cudaStream_t stream, int numSMs, const PartKeys& keys, int64_t rows, uint8_t* pid, unsigned long long* counts);
// pass 2: dst[row] = cursors[pid[row]]++ (cursors = exclusive prefix of counts, device memory)
cudaError_t partitionAssign(cudaStream_t stream, int numSMs, int nparts, int64_t rows, const uint8_t* pid, unsigned long long* cursors, int64_t* dst);
// pass 3 (per column): out[dst[row]] = in[row], element width 1/4/8/16/24/32 bytes
cudaError_t partitionScatter(cudaStream_t stream, int numSMs, int64_t rows, const int64_t* dst, const void* in, void* out, int width);

// ---- gather exchange (small results): push this rank's partial page into its inbox slot on every GPU -----------------------
// Stands in for PartitionedOutputOperator -> OutputBuffer -> ExchangeClient for a partial aggregation whose result is a handful
// of rows (Q1: 4 rows, Q6: 1): no host round trip, no collective. One block per destination copies the first *count rows of
// every output column into the slot (peer memory over NVLink) and then publishes {seq, rows, error bits} with st.release.sys;
// the destination's FINAL kernel (pgpu_k_gather, device/pgpu_device.cuh) waits on that header.
constexpr int kMaxPushCols = 64;
constexpr int kMaxRanks = 64;
struct PushCol {
    const void* values;
    const uint8_t* nulls;
    uint32_t stride;   // bytes per position
    uint32_t pad;
    uint64_t dstV, dstN;  // byte offsets of the column's arrays inside a slot
};
struct PushArgs {
    PushCol c[kMaxPushCols];
    char* slot[kMaxRanks];       // destination r: this rank's slot in r's inbox (header at +0)
    const unsigned long long* count;  // rows in the output buffers (device)
    const uint32_t* err;              // the partial pipeline's sticky error word (device)
    unsigned long long seq;
    int32_t ncols, nranks, capRows, pad;
};
cudaError_t pushPartial(cudaStream_t stream, const PushArgs& args);

// host restatement of the same partition function (tests, multi-rank host logic)
int32_t partitionOfHost(const int64_t* keyBits, const uint8_t* isNull, const int32_t* kinds, int nkeys, int nparts);

}  // namespace pgpu
