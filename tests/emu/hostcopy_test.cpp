// Test driver for presto_b200/csrc/hostcopy.cpp (pure host code): random batches of byte and offset-rebasing tasks, from one
// and from several caller threads at once, checked against plain loops. Built by tests/test_hostcopy.py (also under
// ThreadSanitizer). Test infrastructure only.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <thread>
#include <vector>

#include "../../presto_b200/csrc/hostcopy.h"

using namespace pgpu;

static int runCaller(unsigned seed, int batches, size_t maxBytes) {
    std::mt19937_64 rng(seed);
    int bad = 0;
    for (int b = 0; b < batches; b++) {
        const int ntasks = 1 + (int)(rng() % 9);
        std::vector<std::vector<char>> src(ntasks), dst(ntasks), want(ntasks);
        std::vector<CopyTask> tasks;
        for (int t = 0; t < ntasks; t++) {
            size_t n = (size_t)(rng() % maxBytes);
            const bool offs = rng() % 3 == 0;
            if (offs) n = n / 4 * 4;
            const size_t sa = rng() % 7, da = offs ? 4 * (rng() % 5) : rng() % 67;  // unaligned sources, oddly aligned destinations
            src[t].resize(n + sa + 8);
            dst[t].assign(n + da + 72, (char)0x5a);
            want[t] = dst[t];
            for (size_t i = 0; i < src[t].size(); i++) src[t][i] = (char)rng();
            CopyTask k;
            k.kind = offs ? CopyTask::OFFSETS32 : CopyTask::BYTES;
            k.src = src[t].data() + sa;
            k.dst = dst[t].data() + da;
            k.bytes = n;
            k.delta = offs ? (int32_t)(rng() % 100000) - 50000 : 0;
            tasks.push_back(k);
            if (offs) {
                for (size_t i = 0; i < n / 4; i++) {
                    int32_t v;
                    memcpy(&v, src[t].data() + sa + 4 * i, 4);
                    v = (int32_t)((uint32_t)v + (uint32_t)k.delta);
                    memcpy(want[t].data() + da + 4 * i, &v, 4);
                }
            } else {
                memcpy(want[t].data() + da, src[t].data() + sa, n);
            }
        }
        hostCopy(tasks);
        for (int t = 0; t < ntasks; t++)
            if (dst[t] != want[t]) bad++;
    }
    return bad;
}

int main(int argc, char** argv) {
    const int callers = argc > 1 ? atoi(argv[1]) : 1;
    const int batches = argc > 2 ? atoi(argv[2]) : 50;
    const size_t maxBytes = argc > 3 ? (size_t)atol(argv[3]) : (3u << 20);
    std::vector<int> bad((size_t)callers, 0);
    std::vector<std::thread> th;
    for (int c = 0; c < callers; c++) th.emplace_back([&, c] { bad[c] = runCaller(1234u + 77u * (unsigned)c, batches, maxBytes); });
    for (auto& t : th) t.join();
    int total = 0;
    for (int b : bad) total += b;
    printf("threads=%d callers=%d batches=%d mismatches=%d\n", hostCopyThreads(), callers, batches, total);
    return total ? 1 : 0;
}
