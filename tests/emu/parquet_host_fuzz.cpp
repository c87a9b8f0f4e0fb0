// TEST INFRASTRUCTURE: the host parser of the Parquet ingest (presto_b200/csrc/parquet_host.h) under AddressSanitizer and
// UBSan on mutated chunks. usage: parquet_host_fuzz <chunk file> <physical type> <max def> <num values> <type length> <trials> <seed>
// Every trial flips 1-4 bytes and/or truncates the chunk and parses it with 1 and with 4 page threads; the parser may accept or
// reject, it must not touch memory outside the chunk (the sanitizers abort the process if it does). Prints accepted / rejected.
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iterator>
#include <random>
#include <vector>

#include "../../presto_b200/csrc/parquet_host.h"

int main(int argc, char** argv) {
    if (argc != 8) return 2;
    std::ifstream f(argv[1], std::ios::binary);
    std::vector<uint8_t> raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    const int ptype = atoi(argv[2]), maxDef = atoi(argv[3]), tl = atoi(argv[5]), trials = atoi(argv[6]);
    const long long n = atoll(argv[4]);
    std::mt19937_64 rng((unsigned long long)atoll(argv[7]));
    long accepted = 0, rejected = 0;
    for (int t = 0; t < trials; t++) {
        // an exactly sized heap copy, so that a read one byte past the end is caught
        size_t len = raw.size();
        if (t % 3 == 0 && len) len = (size_t)(rng() % (len + 1));
        std::vector<uint8_t> x(raw.begin(), raw.begin() + (long)len);
        x.shrink_to_fit();
        const int flips = t == 0 ? 0 : 1 + (int)(rng() % 4);
        for (int k = 0; k < flips && len; k++) {
            const size_t span = (t % 2) ? std::min<size_t>(len, 96) : len;   // half of the trials hit the first page header
            x[rng() % span] = (uint8_t)rng();
        }
        for (int threads : {1, 4}) {
            pgpu::pq::Chunk c;
            std::string why;
            if (pgpu::pq::parseChunk(x.data(), x.size(), ptype, maxDef, n, c, why, tl, threads)) {
                accepted++;
                if (ptype == pgpu::pq::BYTE_ARRAY && c.hasDictionary) {
                    std::vector<int32_t> off;
                    std::vector<uint8_t> bytes;
                    pgpu::pq::binaryDictionary(x.data(), c.dictionaryOff, c.dictionaryEnd, c.dictionaryCount, 1 + (int)(rng() % 3), off, bytes, why);
                }
                // the tables must stay inside the chunk: what the device would read
                for (const auto& r : c.valueRuns)
                    if ((r.widthPacked & pgpu::pq::kPackedFlag) && (r.src + (uint64_t)r.count * (r.widthPacked & 0xff) + 7) / 8 > x.size()) return 3;
                for (const auto& r : c.levelRuns)
                    if ((r.widthPacked & pgpu::pq::kPackedFlag) && (r.src + (uint64_t)r.count * (r.widthPacked & 0xff) + 7) / 8 > x.size()) return 4;
                for (const auto& pg : c.pages)
                    if (pg.valuesOff > x.size() || pg.valuesEnd > x.size()) return 5;
            } else {
                rejected++;
                if (why.empty()) return 6;
            }
        }
    }
    printf("%ld %ld\n", accepted, rejected);
    return 0;
}
