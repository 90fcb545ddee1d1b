// Tuning / debugging aid: checks the block directory guess (locate_guess) against the exact block for every step-th BWT position of an .ri file.
#include <cstdlib>
#include "../pangenome-index_b200/csrc/pgi_flatten.hpp"
#include "../pangenome-index_b200/csrc/pgi_core.cuh"
#include <fstream>
#include <iterator>
#include <cstdio>
using namespace pgi;
int main(int argc, char **argv) {
    std::ifstream f(argv[1], std::ios::binary); std::vector<char> b((std::istreambuf_iterator<char>(f)), {});
    RiImage im; flatten_ri(b.data(), b.size(), im);
    RiDev d; d.blocks = im.blocks.data(); d.dir = im.dir.data(); d.sb = im.sb.data(); d.dense = im.dense.data(); d.n = im.k.n; d.n_blocks = im.k.n_blocks; d.dir_shift = im.k.dir_shift; d.sb_shift = im.k.sb_shift; d.has_quirk = 0; d.q_start = d.q_val = nullptr; d.n_quirk = 0;
    uint64_t bad = 0, badleft = 0;
    const uint64_t step = argc > 2 ? strtoull(argv[2], 0, 10) : 1;
    for (uint64_t p = 0; p <= d.n; p += step) {
        uint64_t g = locate_guess(d, p, d.dir[p >> d.dir_shift], d.sb[p >> d.sb_shift]), e = locate_block(d, p);
        if (g != e) ++bad;
        uint32_t w[16]; ld_block(d.blocks + g * 8, w);
        PosRank r = pos_rank(w, p, 8u << 4);
        if (r.left) { if (badleft < 5) printf("p=%llu g=%llu e=%llu left=%u start=%llu E11=%u\n", (unsigned long long) p, (unsigned long long) g, (unsigned long long) e, r.left, (unsigned long long) block_start(w), w[15] >> 16); ++badleft; }
    }
    printf("dense records %llu\n", (unsigned long long) (im.dense.size() / PGI_DENSE_U32));
    printf("n=%llu blocks=%llu dir_shift=%u guess!=exact %llu left!=0 %llu\n", (unsigned long long) d.n, (unsigned long long) d.n_blocks, d.dir_shift, (unsigned long long) bad, (unsigned long long) badleft);
}
