// jit.cpp — see jit.h.
#include "jit.h"

#include <dlfcn.h>
#include <nvrtc.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <mutex>

#include "../../include/prestogpu.h"
#include "plan.h"

// generated at build time from device/pgpu_device.cuh (build.py): const unsigned char pgpu_device_src[]; size_t pgpu_device_src_len
extern "C" const unsigned char pgpu_device_src[];
extern "C" const size_t pgpu_device_src_len;

namespace pgpu {

uint64_t fnv1a64(const void* data, size_t n, uint64_t h) {
    const unsigned char* p = (const unsigned char*)data;
    for (size_t i = 0; i < n; i++) {
        h ^= p[i];
        h *= 0x100000001b3ull;
    }
    return h;
}

std::string jitCacheDir() {
    if (const char* e = getenv("PGPU_CACHE_DIR")) return e;
    Dl_info info;
    if (dladdr((const void*)&jitCacheDir, &info) && info.dli_fname) {
        std::string p = info.dli_fname;
        size_t k = p.rfind('/');
        if (k != std::string::npos) return p.substr(0, k) + "/_kernel_cache";
    }
    return "/tmp/pgpu_kernel_cache";
}

static const char* kOptions[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "--fmad=false", "-device-int128", "-default-device"};

JitResult jitCompile(const std::string& source) {
    JitResult r;
    // key = device library + generated source + options
    uint64_t h = fnv1a64(pgpu_device_src, pgpu_device_src_len);
    h = fnv1a64(source.data(), source.size(), h);
    for (const char* o : kOptions) h = fnv1a64(o, strlen(o), h);
    char name[64];
    snprintf(name, sizeof name, "%016llx.cubin", (unsigned long long)h);
    std::string dir = jitCacheDir();
    std::string path = dir + "/" + name;
    {
        std::ifstream f(path, std::ios::binary);
        if (f) {
            r.cubin.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
            if (r.cubin.size() > 64) {
                r.cacheHit = true;
                return r;
            }
            r.cubin.clear();
        }
    }
    std::string header((const char*)pgpu_device_src, pgpu_device_src_len);
    const char* hdrs[] = {header.c_str()};
    const char* names[] = {"pgpu_device.cuh"};
    nvrtcProgram prog;
    nvrtcResult rc = nvrtcCreateProgram(&prog, source.c_str(), "pgpu_plan.cu", 1, hdrs, names);
    if (rc != NVRTC_SUCCESS) throw PlanError(PGPU_COMPILER_ERROR, std::string("nvrtcCreateProgram: ") + nvrtcGetErrorString(rc));
    rc = nvrtcCompileProgram(prog, (int)(sizeof kOptions / sizeof kOptions[0]), kOptions);
    size_t ls = 0;
    nvrtcGetProgramLogSize(prog, &ls);
    if (ls > 1) {
        r.log.resize(ls);
        nvrtcGetProgramLog(prog, &r.log[0]);
    }
    if (rc != NVRTC_SUCCESS) {
        nvrtcDestroyProgram(&prog);
        if (getenv("PGPU_DUMP_FAILED_SOURCE")) {
            std::ofstream f("/tmp/pgpu_failed_source.cu");
            f << source;
        }
        throw PlanError(PGPU_COMPILER_ERROR, "kernel compilation failed: " + r.log);
    }
    size_t cs = 0;
    nvrtcGetCUBINSize(prog, &cs);
    r.cubin.resize(cs);
    nvrtcGetCUBIN(prog, r.cubin.data());
    nvrtcDestroyProgram(&prog);
    // best-effort cache write (atomic rename)
    mkdir(dir.c_str(), 0755);
    std::string tmp = path + ".tmp" + std::to_string((long)getpid());
    {
        std::ofstream f(tmp, std::ios::binary);
        if (f) {
            f.write(r.cubin.data(), (std::streamsize)r.cubin.size());
            f.close();
            if (rename(tmp.c_str(), path.c_str()) != 0) unlink(tmp.c_str());
        }
    }
    if (getenv("PGPU_KEEP_SOURCE")) {
        std::ofstream f(path.substr(0, path.size() - 6) + ".cu");
        f << source;
    }
    return r;
}

}  // namespace pgpu
