// hostcopy.cpp — see hostcopy.h.
#include "hostcopy.h"

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>

#if defined(__SSE2__)
#include <emmintrin.h>
#define PGPU_HAVE_SSE2 1
#else
#define PGPU_HAVE_SSE2 0
#endif

namespace pgpu {
namespace {

constexpr size_t kSegmentBytes = 256u << 10;  // largest unit of work sharing
constexpr size_t kMinSegmentBytes = 16u << 10;
constexpr size_t kInlineBytes = 48u << 10;    // batches below this are copied by the caller alone

inline void cpuRelax() {
#if PGPU_HAVE_SSE2
    _mm_pause();
#endif
}

// dst gets written once and is read next by the DMA engine: stream it past the caches (no read-for-ownership of the destination
// lines, so the copy costs 2 bytes of DRAM traffic per byte instead of 3)
void copyStream(char* dst, const char* src, size_t n) {
#if PGPU_HAVE_SSE2
    if (n >= 4096) {
        const size_t head = (64 - ((uintptr_t)dst & 63)) & 63;
        memcpy(dst, src, head);
        dst += head;
        src += head;
        n -= head;
        const size_t body = n & ~(size_t)63;
        for (size_t i = 0; i < body; i += 64) {
            const __m128i a = _mm_loadu_si128((const __m128i*)(src + i));
            const __m128i b = _mm_loadu_si128((const __m128i*)(src + i + 16));
            const __m128i c = _mm_loadu_si128((const __m128i*)(src + i + 32));
            const __m128i d = _mm_loadu_si128((const __m128i*)(src + i + 48));
            _mm_stream_si128((__m128i*)(dst + i), a);
            _mm_stream_si128((__m128i*)(dst + i + 16), b);
            _mm_stream_si128((__m128i*)(dst + i + 32), c);
            _mm_stream_si128((__m128i*)(dst + i + 48), d);
        }
        dst += body;
        src += body;
        n -= body;
    }
#endif
    memcpy(dst, src, n);
}

void copyOffsets(int32_t* dst, const char* src, size_t count, int32_t delta) {
    for (size_t i = 0; i < count; i++) {
        int32_t v;
        memcpy(&v, src + 4 * i, 4);  // the source may be unaligned (offsets inside a serialized page)
        dst[i] = v + delta;
    }
}

struct Segment {
    int32_t kind;
    int32_t delta;
    char* dst;
    const char* src;
    size_t bytes;
};

inline int32_t loadI32(const char* p) {
    int32_t v;
    memcpy(&v, p, 4);
    return v;
}

// src[i] >= src[i - 1] for every i of the segment; the element in front of the segment is src[-1] when `hasPred`, else 0
bool ascending32(const char* src, size_t count, bool hasPred) {
    if (!count) return true;
    bool bad = loadI32(src) < (hasPred ? loadI32(src - 4) : 0);
    size_t i = 1;
#if PGPU_HAVE_SSE2
    __m128i acc = _mm_setzero_si128();
    for (; i + 4 <= count; i += 4) {
        const __m128i cur = _mm_loadu_si128((const __m128i*)(src + 4 * i)), prv = _mm_loadu_si128((const __m128i*)(src + 4 * i - 4));
        acc = _mm_or_si128(acc, _mm_cmplt_epi32(cur, prv));
    }
    bad |= _mm_movemask_epi8(acc) != 0;
#endif
    for (; i < count; i++) bad |= loadI32(src + 4 * i) < loadI32(src + 4 * i - 4);
    return !bad;
}

// 0 <= src[i] < limit
bool idsInRange32(const char* src, size_t count, int32_t limit) {
    bool bad = false;
    size_t i = 0;
#if PGPU_HAVE_SSE2
    __m128i acc = _mm_setzero_si128();
    const __m128i zero = _mm_setzero_si128(), top = _mm_set1_epi32(limit - 1);
    for (; i + 4 <= count; i += 4) {
        const __m128i cur = _mm_loadu_si128((const __m128i*)(src + 4 * i));
        acc = _mm_or_si128(acc, _mm_or_si128(_mm_cmplt_epi32(cur, zero), _mm_cmpgt_epi32(cur, top)));
    }
    bad |= _mm_movemask_epi8(acc) != 0;
#endif
    for (; i < count; i++) {
        const int32_t v = loadI32(src + 4 * i);
        bad |= v < 0 || v >= limit;
    }
    return !bad;
}

void runSegment(const Segment& s) {
    if (s.kind == CopyTask::OFFSETS32) copyOffsets((int32_t*)s.dst, s.src, s.bytes / 4, s.delta);
    else if (s.kind == CopyTask::CHECK_ASCENDING32) { if (!ascending32(s.src, s.bytes / 4, s.delta != 0)) __atomic_store_n((int32_t*)s.dst, 1, __ATOMIC_RELAXED); }
    else if (s.kind == CopyTask::CHECK_IDS32) { if (!idsInRange32(s.src, s.bytes / 4, s.delta)) __atomic_store_n((int32_t*)s.dst, 1, __ATOMIC_RELAXED); }
    else copyStream(s.dst, s.src, s.bytes);
}

void finishStores() {
#if PGPU_HAVE_SSE2
    _mm_sfence();  // non-temporal stores must be globally visible before the copy engine is told to read them
#endif
}

class Pool {
public:
    Pool() {
        int n = 0;
        if (const char* e = getenv("PGPU_COPY_THREADS")) n = atoi(e);
        if (n <= 0) {
            const unsigned hw = std::thread::hardware_concurrency();
            n = (int)std::min(8u, std::max(1u, hw / 2));
        }
        nthreads_ = std::max(1, std::min(n, 64));
        for (int i = 1; i < nthreads_; i++) workers_.emplace_back([this] { workerLoop(); });
    }
    ~Pool() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_.store(true);
            gen_.fetch_add(1);
        }
        cv_.notify_all();
        for (auto& t : workers_) t.join();
    }
    int threads() const { return nthreads_; }

    void run(const std::vector<Segment>& segs) {
        std::unique_lock<std::mutex> own(runMu_, std::try_to_lock);
        if (!own.owns_lock() || workers_.empty()) {  // another driver thread is using the workers: this one copies alone
            for (const Segment& s : segs) runSegment(s);
            finishStores();
            return;
        }
        segs_ = &segs;
        total_ = segs.size();
        next_.store(0);
        finished_.store(0);
        open_.store(true);
        gen_.fetch_add(1);
        if (sleepers_.load() > 0) {
            std::lock_guard<std::mutex> lk(m_);
            cv_.notify_all();
        }
        drain();
        while (finished_.load() < total_) cpuRelax();
        open_.store(false);
        while (busy_.load() != 0) cpuRelax();  // no worker still looks at this batch when the caller's vector goes away
    }

private:
    void drain() {
        size_t i;
        bool any = false;
        while ((i = next_.fetch_add(1)) < total_) {
            runSegment((*segs_)[i]);
            any = true;
            finished_.fetch_add(1);
        }
        if (any) finishStores();
    }
    void workerLoop() {
        uint64_t seen = gen_.load();
        for (;;) {
            int spins = 0;
            while (gen_.load() == seen && !stop_.load()) {
                if (++spins < 20000) {
                    cpuRelax();
                } else {
                    std::unique_lock<std::mutex> lk(m_);
                    sleepers_.fetch_add(1);
                    cv_.wait(lk, [&] { return gen_.load() != seen || stop_.load(); });
                    sleepers_.fetch_sub(1);
                }
            }
            if (stop_.load()) return;
            seen = gen_.load();
            busy_.fetch_add(1);
            if (open_.load()) {  // the batch fields are stable for as long as busy_ is held
                size_t i;
                while ((i = next_.fetch_add(1)) < total_) {
                    runSegment((*segs_)[i]);
                    finishStores();  // before the segment is reported done: the caller starts the DMA right after
                    finished_.fetch_add(1);
                }
            }
            busy_.fetch_sub(1);
        }
    }

    int nthreads_ = 1;
    std::vector<std::thread> workers_;
    std::mutex runMu_, m_;
    std::condition_variable cv_;
    std::atomic<uint64_t> gen_{0};
    std::atomic<bool> stop_{false}, open_{false};
    std::atomic<int> sleepers_{0}, busy_{0};
    const std::vector<Segment>* segs_ = nullptr;
    size_t total_ = 0;
    std::atomic<size_t> next_{0}, finished_{0};
};

Pool& pool() {
    static Pool* p = new Pool();  // intentionally leaked: worker threads must not be joined from a static destructor at exit
    return *p;
}

}  // namespace

int hostCopyThreads() { return pool().threads(); }

void hostCopy(const std::vector<CopyTask>& tasks) {
    size_t total = 0;
    for (const CopyTask& t : tasks) total += t.bytes;
    if (total == 0) return;
    // a 4096-row page (RecordPageSource's shape) is ~190 KB: cut small batches finer so that every worker gets a share
    size_t seg = total / (size_t)(2 * hostCopyThreads());
    seg = std::min(kSegmentBytes, std::max(kMinSegmentBytes, (seg + 4095) / 4096 * 4096));
    std::vector<Segment> segs;
    segs.reserve(tasks.size() + total / seg + 1);
    for (const CopyTask& t : tasks) {
        for (size_t at = 0; at < t.bytes; at += seg) {
            const size_t n = std::min(seg, t.bytes - at);
            if (t.kind == CopyTask::CHECK_ASCENDING32) segs.push_back(Segment{(int32_t)t.kind, at ? 1 : 0, (char*)t.dst, (const char*)t.src + at, n});  // delta: has a predecessor
            else if (t.kind == CopyTask::CHECK_IDS32) segs.push_back(Segment{(int32_t)t.kind, t.delta, (char*)t.dst, (const char*)t.src + at, n});
            else segs.push_back(Segment{(int32_t)t.kind, t.delta, (char*)t.dst + at, (const char*)t.src + at, n});
        }
    }
    if (total < kInlineBytes) {
        for (const Segment& s : segs) runSegment(s);
        finishStores();
        return;
    }
    pool().run(segs);
}

}  // namespace pgpu
