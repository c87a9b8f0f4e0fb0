// codec_test.cpp — the host-side page decoders of presto_b200/csrc/parquet_host.h (Snappy, LZ4 block, LZ4 Hadoop frames, gzip /
// DEFLATE, Zstandard) on files: codec_test <codec> <compressed file> <uncompressed size> <output file>; exit 0 = decoded (output
// written), 3 = the decoder refused the input. Built with -fsanitize=address,undefined by tests/test_codecs.py: damaged inputs
// must be refused or decode to garbage, never read or write out of bounds.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "parquet_host.h"

int main(int argc, char** argv) {
    if (argc != 5) return 2;
    const std::string codec = argv[1];
    FILE* f = fopen(argv[2], "rb");
    if (!f) return 2;
    std::vector<uint8_t> in;
    uint8_t buf[65536];
    size_t got;
    while ((got = fread(buf, 1, sizeof buf, f)) > 0) in.insert(in.end(), buf, buf + got);
    fclose(f);
    const size_t outLen = (size_t)atoll(argv[3]);
    // exact-size heap buffers: the sanitizer sees every byte outside them
    uint8_t* src = (uint8_t*)malloc(in.size() ? in.size() : 1);
    memcpy(src, in.data(), in.size());
    uint8_t* out = (uint8_t*)malloc(outLen ? outLen : 1);
    using namespace pgpu::pq;
    bool ok;
    if (codec == "snappy") ok = snappyDecompress(src, in.size(), out, outLen);
    else if (codec == "lz4_raw") ok = lz4Decompress(src, in.size(), out, outLen);
    else if (codec == "lz4_hadoop") ok = lz4HadoopDecompress(src, in.size(), out, outLen);
    else if (codec == "gzip") ok = gzipDecompress(src, in.size(), out, outLen);
    else if (codec == "zstd") ok = pgpu::zstd::decompress(src, in.size(), out, outLen);
    else return 2;
    if (ok) {
        FILE* o = fopen(argv[4], "wb");
        if (!o) return 2;
        fwrite(out, 1, outLen, o);
        fclose(o);
    }
    free(src);
    free(out);
    return ok ? 0 : 3;
}
