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

// CHECK_ASCENDING32 / CHECK_IDS32: one planted fault anywhere in a (possibly unaligned, multi-segment) array must be found,
// a clean array must pass
static int runChecks(unsigned seed) {
    std::mt19937 rng(seed);
    int bad = 0;
    for (int trial = 0; trial < 200; trial++) {
        const size_t n = 1 + rng() % 700000;
        const size_t pad = rng() % 4;  // unaligned start, as inside a serialized page
        std::vector<char> buf(pad + 4 * n);
        std::vector<int32_t> v(n);
        int32_t at = 0;
        for (size_t i = 0; i < n; i++) v[i] = (at += (int32_t)(rng() % 3));
        const bool plant = trial % 2 == 1;
        if (plant) {
            const size_t where = rng() % n;
            v[where] = where ? v[where - 1] - 1 - (int32_t)(rng() % 5) : -1;
        }
        memcpy(buf.data() + pad, v.data(), 4 * n);
        int32_t flag = 0;
        hostCopy({CopyTask{CopyTask::CHECK_ASCENDING32, &flag, buf.data() + pad, 4 * n, 0}});
        if ((flag != 0) != plant) bad++;
        const int32_t limit = 1 + (int32_t)(rng() % 100000);
        for (size_t i = 0; i < n; i++) v[i] = (int32_t)(rng() % (uint32_t)limit);
        if (plant) v[rng() % n] = trial % 4 == 1 ? limit + (int32_t)(rng() % 3) : -1 - (int32_t)(rng() % 3);
        memcpy(buf.data() + pad, v.data(), 4 * n);
        flag = 0;
        hostCopy({CopyTask{CopyTask::CHECK_IDS32, &flag, buf.data() + pad, 4 * n, limit}});
        if ((flag != 0) != plant) bad++;
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
    int total = runChecks(99u);
    for (int b : bad) total += b;
    printf("threads=%d callers=%d batches=%d mismatches=%d\n", hostCopyThreads(), callers, batches, total);
    return total ? 1 : 0;
}
