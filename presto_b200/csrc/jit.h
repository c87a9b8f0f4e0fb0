// jit.h — NVRTC compilation of generated plan kernels to sm_100a cubins, with an on-disk cache.
// Stands where the reference has its runtime bytecode pipeline: main/sql/gen/ExpressionCompiler.java:74-137
// (cache + compile) and airlift bytecode's DynamicClassLoader.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace pgpu {

struct JitResult {
    std::vector<char> cubin;
    bool cacheHit = false;
    std::string log;
};

// Compiles `source` (which #includes "pgpu_device.cuh") for sm_100a. Throws PlanError(PGPU_COMPILER_ERROR).
// Needs no GPU.
JitResult jitCompile(const std::string& source);

// Directory of the cubin cache (PGPU_CACHE_DIR, else <dir of libprestogpu.so>/_kernel_cache).
std::string jitCacheDir();

uint64_t fnv1a64(const void* data, size_t n, uint64_t seed = 0xcbf29ce484222325ull);

}  // namespace pgpu
