// TEST INFRASTRUCTURE: plan parser + CUDA code generator (presto_b200/csrc/plan.cpp, codegen.cpp) under AddressSanitizer + UBSan.
// usage: codegen_asan <plan file>...   For every plan: generate the kernel source for every combination of block encodings
// (plain / dictionary / RLE, up to 81 combinations; the rest sampled) with and without null vectors. Unsupported combinations may
// throw PlanError; the generator must not misbehave. Prints the number of sources generated and their total size.
#include <cstdio>
#include <fstream>
#include <iterator>
#include <string>
#include <vector>

#include "../../presto_b200/csrc/codegen.h"

int main(int argc, char** argv) {
    long generated = 0, rejected = 0;
    size_t bytes = 0;
    for (int a = 1; a < argc; a++) {
        std::ifstream f(argv[a]);
        const std::string text((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
        std::unique_ptr<pgpu::Plan> plan;
        try {
            plan = pgpu::parsePlan(text);
        } catch (const pgpu::PlanError&) {
            return 3;  // the plans handed in are valid
        }
        const size_t nc = plan->input.size();
        size_t combos = 1;
        for (size_t c = 0; c < nc && combos < 81; c++) combos *= 3;
        for (size_t k = 0; k < combos; k++) {
            for (int nullable = 0; nullable < 2; nullable++) {
                std::vector<pgpu::ColSig> sig(nc);
                size_t v = k;
                for (size_t c = 0; c < nc; c++) {
                    sig[c].encoding = (int)(v % 3);
                    v = v / 3 + (c * 7 + k) % 3 * (nc > 4);  // beyond 4 channels: a deterministic sample
                    sig[c].nullable = nullable != 0;
                }
                try {
                    pgpu::GenResult g = pgpu::generate(*plan, sig);
                    generated++;
                    bytes += g.source.size();
                    if (g.source.find("pgpu_k_page") == std::string::npos) return 4;
                } catch (const pgpu::PlanError&) {
                    rejected++;
                }
            }
        }
    }
    printf("%ld %ld %zu\n", generated, rejected, bytes);
    return 0;
}
