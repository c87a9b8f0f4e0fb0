// parquet_ingest.h — one Parquet row group -> the device Blocks of one input Page (see parquet_host.h / parquet.h for the two halves
// it puts together). Kept apart from the pipeline so that the same code is driven by pgpu_add_parquet_row_group on the GPU and, with
// the CUDA calls and kernel launches replaced, by the CPU emulation of tests/emu/parquet_ingest_emu.cpp.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <functional>
#include <string>
#include <vector>

#include "../../include/prestogpu.h"
#include "plan.h"

namespace pgpu {

struct PqDecoded {
    std::vector<pgpu_column> cols, dicts;  // cols[c].dictionary points into dicts
    int64_t h2dBytes = 0;
    int launches = 0;
};

// `arena(bytes)` returns device memory of at least that size that stays valid until the consumer of the columns has run on `stream`
// (nullptr: out of memory). `dErr` = the pipeline's sticky device error word. On failure `why` holds the message.
pgpu_status pqDecodeRowGroup(const Plan& plan, const pgpu_parquet_chunk* chunks, int32_t nchunks, int64_t rows, cudaStream_t stream, uint32_t* dErr, int pageThreads,
                             const std::function<char*(size_t)>& arena, PqDecoded& out, std::string& why);

}  // namespace pgpu
