// build_rindex <x.rl_bwt> [--legacy] [--no-samples] > x.ri
//
// Replaces src/build_rindex.cpp:11-20 of the reference: FastLocate(rlbwt_file) (src/r-index.cpp:778-1139) followed by
// serialize_encoded(std::cout) (src/r-index.cpp:297-376). --legacy writes FastLocate::serialize instead (the layout of
// the shipped test_data/*.ri files, which this tool regenerates byte for byte). CPU data preparation around the hot
// path: the files it writes load in the reference and in pgi_index_create alike.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iterator>
#include <string>
#include <vector>

extern "C" {
const char *prep_last_error();
void prep_free(void *p);
int prep_rindex_from_rlbwt(const uint8_t *rlbwt, uint64_t len, int encoded, int with_samples,
                           uint8_t **ri, uint64_t *ri_len, uint64_t *row_textpos, uint64_t *info);
}

int main(int argc, char **argv) {
    if (argc < 2) { std::fprintf(stderr, "Usage: %s <x.rl_bwt> [--legacy] [--no-samples] > x.ri\n", argv[0]); return EXIT_FAILURE; }
    int encoded = 1, samples = 1;
    for (int i = 2; i < argc; ++i) {
        if (!std::strcmp(argv[i], "--legacy")) encoded = 0;
        else if (!std::strcmp(argv[i], "--no-samples")) samples = 0;
        else { std::fprintf(stderr, "build_rindex: unknown option %s\n", argv[i]); return EXIT_FAILURE; }
    }
    std::ifstream in(argv[1], std::ios::binary);
    if (!in) { std::fprintf(stderr, "build_rindex: cannot open %s\n", argv[1]); return EXIT_FAILURE; }
    std::vector<char> rl((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
    uint8_t *ri = nullptr; uint64_t len = 0, info[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (prep_rindex_from_rlbwt((const uint8_t *) rl.data(), rl.size(), encoded, samples, &ri, &len, nullptr, info) != 0) {
        std::fprintf(stderr, "build_rindex: %s\n", prep_last_error()); return EXIT_FAILURE;
    }
    std::fprintf(stderr, "build_rindex: n = %llu, sequences = %llu, max_length = %llu, runs = %llu, sigma = %llu, %s layout, %llu bytes\n",
                 (unsigned long long) info[0], (unsigned long long) info[1], (unsigned long long) info[2], (unsigned long long) info[3],
                 (unsigned long long) info[4], encoded ? "ENCODED_BLOCKS" : "legacy", (unsigned long long) len);
    const bool ok = std::fwrite(ri, 1, len, stdout) == len && std::fflush(stdout) == 0;
    prep_free(ri);
    return ok ? 0 : EXIT_FAILURE;
}
