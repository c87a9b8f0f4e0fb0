// hostcopy.h — the host half of Operator.addInput for pages that live in ordinary (pageable) memory: Java heap arrays behind
// LongArrayBlock.values / VariableWidthBlock.offsets (spi/block/LongArrayBlock.java:37-41) cannot be page-locked by the caller,
// so every byte crosses the CPU once on its way into the pinned staging ring. A single thread copies ~10 GB/s, a fifth of the
// PCIe Gen5 link; this pool splits one page's arrays over worker threads (non-temporal stores: the destination is read next by
// the copy engine, not by a core) so the staging copy keeps up with the DMA it feeds.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace pgpu {

struct CopyTask {
    // CHECK_*: nothing is copied; `dst` points to an int32 the task sets to 1 when the source fails the check (wire-format pages:
    // the device follows these arrays blindly). CHECK_ASCENDING32: src[0] >= 0 and src[i] >= src[i - 1]; CHECK_IDS32: 0 <= src[i] < delta.
    enum Kind : int32_t { BYTES = 0, OFFSETS32 = 1, CHECK_ASCENDING32 = 2, CHECK_IDS32 = 3 };
    Kind kind = BYTES;
    void* dst = nullptr;        // BYTES: destination; OFFSETS32: int32 destination (4-byte aligned)
    const void* src = nullptr;  // may be unaligned (wire-format payloads)
    size_t bytes = 0;           // BYTES: byte count; OFFSETS32: entry count * 4
    int32_t delta = 0;          // OFFSETS32: dst[i] = src[i] + delta (rebasing VariableWidthBlock offsets onto the super-page)
};

// Runs every task to completion before returning (the caller's buffers may be reused afterwards). Large batches are split into
// segments and shared with the pool's workers; the calling thread always takes part. Safe to call from several threads: a caller
// that finds the pool busy copies its own batch alone (that is the multi-driver case, which is parallel by itself).
void hostCopy(const std::vector<CopyTask>& tasks);

// Worker threads the pool uses (PGPU_COPY_THREADS, default min(16, max(1, hardware threads / 2)) incl. the caller).
int hostCopyThreads();

}  // namespace pgpu
