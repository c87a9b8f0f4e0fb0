// tableutil.h — small ahead-of-time kernels around the HBM hash table that do not depend on the plan.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace pgpu {

// Clustering probe for the bucketed table (device/pgpu_device.cuh "bucketed slots"): over the first n (<= 65536) keys of an
// integer key column, out[0] = positions whose key differs from its predecessor's, out[1 + d] (d = 0..7) = those of them whose
// predecessor lies in the same bucket of 8 when buckets are cut at 8n - d ((key + d) >> 3 equal): clusters of 8 keys need not
// start at a multiple of 8 (TPC-H order keys start at 8n + 1). out[9] = positions whose key is smaller than its predecessor's, compared
// as signed integers (the streaming-runs path expects its first grouping field in that order). `width` = 4 or 8 bytes per key. out[0..9] zeroed by the caller.
cudaError_t keyClustering(cudaStream_t stream, const void* keys, int width, int64_t n, unsigned long long* out);

// Partition sweep (device/pgpu_device.cuh "partition sweep"): rows per partition -> first row of every partition.
// cursor[k] = start[k] = hist[0] + ... + hist[k - 1]; start[K] = the total. K <= 1024.
// streaming runs, two passes: offsets[i] = counts[0] + ... + counts[i - 1], *total = the sum of all n counts. `sums`: 1024 words of scratch;
// nblocks: how many blocks share the work (<= 1024).
cudaError_t runsScan(cudaStream_t stream, const unsigned* counts, long long n, unsigned long long* offsets, unsigned long long* sums, int nblocks,
                     unsigned long long* total);
cudaError_t partitionScan(cudaStream_t stream, const unsigned* hist, unsigned long long* cursor, unsigned long long* start, int K);

}  // namespace pgpu
