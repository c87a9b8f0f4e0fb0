// TEST INFRASTRUCTURE: the plan parser (presto_b200/csrc/plan.cpp) under AddressSanitizer + UBSan on mutated plan texts.
// usage: plan_fuzz <plan file> <trials> <seed>. A plan is accepted or rejected with a PlanError; nothing else may happen.
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iterator>
#include <random>
#include <string>

#include "../../presto_b200/csrc/plan.h"

int main(int argc, char** argv) {
    if (argc != 4) return 2;
    std::ifstream f(argv[1]);
    const std::string raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    std::mt19937_64 rng((unsigned long long)atoll(argv[3]));
    const char alphabet[] = "() 0123456789-xabcdefnilousvwhpt";
    long accepted = 0, rejected = 0;
    for (int t = 0; t < atoi(argv[2]); t++) {
        std::string x = raw;
        const int edits = t == 0 ? 0 : 1 + (int)(rng() % 4);
        for (int k = 0; k < edits && !x.empty(); k++) {
            const size_t at = rng() % x.size();
            switch (rng() % 4) {
                case 0: x[at] = alphabet[rng() % (sizeof alphabet - 1)]; break;
                case 1: x.erase(at, 1 + rng() % 6); break;
                case 2: x.insert(at, 1, alphabet[rng() % (sizeof alphabet - 1)]); break;
                default: x.resize(at); break;
            }
        }
        try {
            auto plan = pgpu::parsePlan(x);
            (void)plan->outputTypes();
            accepted++;
        } catch (const pgpu::PlanError& e) {
            rejected++;
            if (!*e.what()) return 3;
        }
    }
    printf("%ld %ld\n", accepted, rejected);
    return 0;
}
