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

// host restatement of the same partition function (tests, multi-rank host logic)
int32_t partitionOfHost(const int64_t* keyBits, const uint8_t* isNull, const int32_t* kinds, int nkeys, int nparts);

}  // namespace pgpu
