// pgi_flatten.hpp — host side of the index: parse the reference's on-disk files and flatten them
// into the device layout (DESIGN.md "Data layout in HBM"). Pure C++17, no CUDA, no SDSL/GBWT.
//
// Replaces (reference file:line):
//   FastLocate::load_encoded / load            src/r-index.cpp:406-459 / :378-404
//   Run_blocks::load                           include/pangenome_index/r-index.hpp:283-292
//   EncodedBlock stream grammar                src/r-index.cpp:65-156
//   TagArray::load_compressed_tags_sdsl        src/tag_arrays.cpp:766-776
//   TagArray::load_compressed_tags             src/tag_arrays.cpp:739-763
// The SDSL containers (int_vector, bit_vector, sd_vector + its two select supports) and
// gbwt::ByteCode are read straight from their serialized form.
//
// Nothing here answers queries on the CPU: the output is a set of flat arrays that pgi_capi.cu
// uploads; all query arithmetic is in pgi_core.cuh and runs on the GPU.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include "pgi_layout.h"

namespace pgi {

struct FormatError : std::runtime_error { using std::runtime_error::runtime_error; };

// ---------------------------------------------------------------------------------- byte reader
class Reader {
public:
    Reader(const void *p, size_t n) : base_((const uint8_t *) p), len_(n), off_(0) {}
    size_t tell() const { return off_; }
    size_t remaining() const { return len_ - off_; }
    const uint8_t *take(size_t n) {
        if (n > len_ - off_) throw FormatError("truncated file at byte " + std::to_string(off_));
        const uint8_t *r = base_ + off_; off_ += n; return r;
    }
    template <typename T> T get() { T v; std::memcpy(&v, take(sizeof(T)), sizeof(T)); return v; }
private:
    const uint8_t *base_; size_t len_, off_;
};

// A serialized sdsl::int_vector viewed in place (payload words are not copied).
struct IntVectorView {
    uint64_t bits = 0; unsigned width = 0; const uint8_t *payload = nullptr;
    // fixed == 0 -> int_vector<0> (width byte present)
    void parse(Reader &r, unsigned fixed) {
        bits = r.get<uint64_t>();
        width = fixed ? fixed : r.get<uint8_t>();
        payload = r.take((size_t) ((bits + 63) / 64) * 8);
    }
    uint64_t size() const { return width ? bits / width : 0; }
    uint64_t word(uint64_t w) const { uint64_t v; std::memcpy(&v, payload + 8 * w, 8); return v; }
    uint64_t at(uint64_t i) const {
        uint64_t b = i * width, w = b >> 6; unsigned s = (unsigned) (b & 63);
        uint64_t v = word(w) >> s;
        if (s + width > 64) v |= word(w + 1) << (64 - s);
        return width >= 64 ? v : v & ((1ULL << width) - 1);
    }
};

inline void skip_select_support(Reader &r) {
    uint64_t cnt = r.get<uint64_t>();
    if (!cnt) return;
    IntVectorView v; v.parse(r, 0); v.parse(r, 1);
    for (uint64_t i = 0, sb = (cnt + 4095) / 4096; i < sb; ++i) v.parse(r, 0);
}

// A serialized sdsl::sd_vector<> viewed in place; enumerates its 1-positions in increasing order.
struct SdVectorView {
    uint64_t size = 0; unsigned wl = 0; IntVectorView low, high; uint64_t ones = 0;
    void parse(Reader &r) {
        size = r.get<uint64_t>(); wl = r.get<uint8_t>();
        low.parse(r, 0); high.parse(r, 1);
        skip_select_support(r); skip_select_support(r);
        ones = 0;
        for (uint64_t w = 0, nw = (high.bits + 63) / 64; w < nw; ++w) {
            uint64_t x = high.word(w);
            if ((w + 1) * 64 > high.bits) x &= (high.bits & 63) ? ((1ULL << (high.bits & 63)) - 1) : ~0ULL;
            ones += (uint64_t) __builtin_popcountll(x);
        }
        if (wl && low.size() != ones) throw FormatError("sd_vector: low/high parts disagree");
    }
    template <typename F> void for_each_one(F &&f) const {
        uint64_t j = 0;
        for (uint64_t w = 0, nw = (high.bits + 63) / 64; w < nw; ++w) {
            uint64_t x = high.word(w);
            if ((w + 1) * 64 > high.bits) x &= (high.bits & 63) ? ((1ULL << (high.bits & 63)) - 1) : ~0ULL;
            while (x) {
                unsigned b = (unsigned) __builtin_ctzll(x); x &= x - 1;
                uint64_t hi = w * 64 + b - j;
                f(j, (hi << wl) | (wl ? low.at(j) : 0));
                ++j;
            }
        }
    }
};

// gbwt::ByteCode value at p (7 data bits per byte, low group first, bit 7 = continue).
inline uint64_t bytecode_next(const uint8_t *&p, const uint8_t *end) {
    uint64_t v = 0; unsigned sh = 0;
    for (;;) {
        if (p >= end) throw FormatError("ByteCode value runs past the end of its stream");
        uint8_t b = *p++;
        if (sh < 64) v |= (uint64_t) (b & 0x7F) << sh;
        sh += 7;
        if (!(b & 0x80)) return v;
    }
}

// ---------------------------------------------------------------------------------- flat images
struct SparseRankImage {
    SparseRankConsts k; std::vector<uint64_t> dir; std::vector<uint8_t> low;
};
// A bit-packed sdsl::int_vector<0> kept exactly as in the file (+ 2 spare words: a 2-word read never leaves the array).
struct PackedImage { std::vector<uint64_t> words; uint64_t n = 0; unsigned width = 0; };
inline void pack_from_view(const IntVectorView &v, PackedImage &im) {
    im.n = v.size(); im.width = v.width;
    const size_t nw = (size_t) ((v.bits + 63) / 64);
    im.words.assign(nw + 2, 0);
    if (nw) std::memcpy(im.words.data(), v.payload, nw * 8);
}
struct RiImage {
    RiConsts k;                          // scalars + tables copied to the device as one struct
    std::vector<uint64_t> blocks;        // PGI_BLOCK_WORDS u64 per 64-byte block: one per bucket, then the chunk blocks
    std::vector<uint64_t> q_start, q_val;// quirk slot tables (legacy files with an absent nuc symbol)
    uint64_t n_file_runs = 0, n_ref_blocks = 0; bool encoded = false;
    // locate (toehold samples + phi), only when the file carries samples (pgi_layout.h: locate_ok)
    SparseRankImage run_starts, last;    // first BWT row of every logical run; `last` (SA values at the last row of every run)
    PackedImage samples, last_to_run;
};
struct TagImage {
    int layout = 0; SparseRankImage intervals;
    std::vector<uint64_t> starts;        // select_1(j + 1) of encoded_runs_starts_sd
    uint64_t n_items = 0; unsigned width = 0;
    std::vector<uint64_t> words;         // sdsl-compact items, bit-packed exactly as in the file
    std::vector<uint8_t> stream;         // ByteCode layout
    bool starts_regular = false;         // starts[j] == 10 j for every j (lookup collapses to arithmetic)
};

static const unsigned char kNuc[6] = {'\n', 'A', 'C', 'G', 'N', 'T'};   // include/pangenome_index/utils.hpp:11

// ---------------------------------------------------------------------------------- run blocks
// Collects the (code, length) runs of the file (adjacent runs of one symbol merged), 4 bytes per run.
class RunCollector {
public:
    std::vector<uint32_t> runs;          // code | length << 3, length < 2^29 (longer runs are split)
    uint64_t lcum[6] = {0, 0, 0, 0, 0, 0}; uint64_t lpos = 0;    // counts / positions of the runs added so far
    void add(unsigned code, uint64_t len) {
        if (!len) return;
        lcum[code] += len; lpos += len;
        if (code == pcode_) { plen_ += len; return; }
        flush(); pcode_ = code; plen_ = len;
    }
    void finish() { flush(); pcode_ = 8; }
private:
    unsigned pcode_ = 8; uint64_t plen_ = 0;
    void flush() {
        const uint64_t M = (1ull << 29) - 1;
        while (plen_) { const uint64_t t = plen_ < M ? plen_ : M; runs.push_back((uint32_t) (pcode_ | t << 3)); plen_ -= t; }
    }
};

// At most one bucket in crowded_div() may be crowded (PGI_CROWDED_DIV=<n> overrides the default for tuning runs).
inline uint64_t crowded_div() {
    if (const char *e = std::getenv("PGI_CROWDED_DIV")) { const long v = std::strtol(e, nullptr, 10); if (v > 0) return (uint64_t) v; }
    return PGI_CROWDED_DIV;
}
// Buckets holding more than PGI_RUNS_PER_BLOCK clipped runs for a bucket size of 2^S positions.
inline uint64_t count_overflow_buckets(const std::vector<uint32_t> &runs, unsigned S) {
    uint64_t pos = 0, cur = 0, cnt = 0, over = 0;      // cur: bucket being filled, cnt: its entries so far
    for (uint32_t e : runs) {
        const uint64_t len = e >> 3, b0 = pos >> S, b1 = (pos + len - 1) >> S;
        if (b0 != cur) { over += cnt > PGI_RUNS_PER_BLOCK; cur = b0; cnt = 0; }
        ++cnt;
        if (b1 != b0) { over += cnt > PGI_RUNS_PER_BLOCK; cur = b1; cnt = 1; }   // buckets strictly between hold one entry
        pos += len;
    }
    return over + (cnt > PGI_RUNS_PER_BLOCK);
}

// One 64-byte block (pgi_layout.h) from up to 12 entries; cum = symbol counts before the first entry.
inline void write_block(uint64_t *dst, const uint64_t cum[6], const uint32_t *code, const uint32_t *end, unsigned ne, uint32_t w0) {
    if (cum[1] >> 40 || cum[2] >> 40 || cum[3] >> 40 || cum[5] >> 40)
        throw FormatError("index longer than 2^40 positions is not supported by the device layout");
    if (cum[0] >> 32 || cum[4] >> 32)
        throw FormatError("more than 2^32 endmarkers or N symbols is not supported by the device layout");
    uint32_t w[PGI_BLOCK_U32];
    w[0] = w0;
    uint64_t codes = 0;
    for (unsigned i = 0; i < PGI_RUNS_PER_BLOCK; ++i) codes |= (uint64_t) (i < ne ? code[i] : PGI_CODE_UNUSED) << (4 * i);
    auto e = [&](unsigned i) -> uint32_t { return ne ? end[i < ne ? i : ne - 1] : 0u; };     // unused runs repeat the last end offset
    w[8] = (uint32_t) (codes >> 32) << 16;
    pgi_put_header(w, cum);
    w[9] = (uint32_t) codes;
    for (unsigned j = 0; j < 6; ++j) w[10 + j] = e(2 * j) | e(2 * j + 1) << 16;
    for (unsigned j = 0; j < PGI_BLOCK_WORDS; ++j) dst[j] = (uint64_t) w[2 * j] | (uint64_t) w[2 * j + 1] << 32;
}

// Crowded bucket (more than 12 clipped runs, given as code[] / end[] with end offsets relative to the bucket): the largest
// sub_shift for which every sub-bucket of 2^sub_shift positions holds at most PGI_CHUNK_RUNS clipped runs.
inline unsigned choose_sub_shift(const std::vector<uint32_t> &end, unsigned S) {
    for (unsigned ss = S - 1;; --ss) {
        bool ok = true; size_t i = 0; const uint32_t n_sub = 1u << (S - ss);
        for (uint32_t c = 0; c < n_sub && ok; ++c) {
            const uint32_t hi = (c + 1) << ss; unsigned k = 0;
            while (i < end.size()) { ++k; if (end[i] > hi) break; ++i; if (end[i - 1] == hi) break; }
            ok = k <= PGI_CHUNK_RUNS;
        }
        if (ok || ss == 0) return ss;
    }
}

// The bucket array (pgi_layout.h): block b = positions [b << S, (b + 1) << S); chunk blocks of crowded buckets behind.
// totals[c] receives the number of occurrences of code c.
inline void build_buckets(RiImage &im, const std::vector<uint32_t> &runs, uint64_t n, uint64_t totals[6]) {
    if (runs.empty()) throw FormatError("r-index has no runs");
    unsigned S = PGI_MAX_BSHIFT;
    if (const char *e = std::getenv("PGI_BSHIFT")) {
        S = (unsigned) std::atoi(e);
        if (S < PGI_MIN_BSHIFT || S > PGI_MAX_BSHIFT) throw FormatError("PGI_BSHIFT outside [4, 12]");
    } else {
        for (; S > PGI_MIN_BSHIFT; --S) if (count_overflow_buckets(runs, S) <= ((n >> S) + 1) / crowded_div()) break;
    }
    const uint64_t nb = (n >> S) + 1, bsize = 1ull << S;
    im.k.bshift = S; im.k.n_buckets = nb;
    im.blocks.assign((size_t) nb * PGI_BLOCK_WORDS, 0);
    uint64_t cum[6] = {0, 0, 0, 0, 0, 0}, pos = 0, n_dir = 0;
    size_t ri = 0; uint64_t carry_len = 0; unsigned carry_code = 0;        // part of a run that continues into this bucket
    std::vector<uint32_t> code, end;
    for (uint64_t b = 0; b < nb; ++b) {
        const uint64_t lo = b << S, hi = lo + bsize;
        uint64_t bcum[6]; std::memcpy(bcum, cum, sizeof(cum));
        code.clear(); end.clear();
        while (pos < hi && (carry_len || ri < runs.size())) {
            if (!carry_len) { carry_code = runs[ri] & 7u; carry_len = runs[ri] >> 3; ++ri; }
            const uint64_t t = carry_len < hi - pos ? carry_len : hi - pos;
            if (!code.empty() && code.back() == carry_code) end.back() += (uint32_t) t;     // a run split at 2^29 by the collector
            else { code.push_back(carry_code); end.push_back((uint32_t) (pos + t - lo)); }
            cum[carry_code] += t; pos += t; carry_len -= t;
        }
        uint64_t *dst = im.blocks.data() + (size_t) b * PGI_BLOCK_WORDS;
        const unsigned ne = (unsigned) code.size();
        if (ne <= PGI_RUNS_PER_BLOCK) { write_block(dst, bcum, code.data(), end.data(), ne, 0u); continue; }
        // crowded bucket: directory block + one chunk block per sub-bucket of 2^ss positions (pad entry + at most 11 runs)
        const unsigned ss = choose_sub_shift(end, S); const uint32_t n_sub = 1u << (S - ss);
        const uint64_t base = im.blocks.size() / PGI_BLOCK_WORDS;
        im.blocks.resize(im.blocks.size() + (size_t) n_sub * PGI_BLOCK_WORDS);
        uint64_t ccum[6]; std::memcpy(ccum, bcum, sizeof(bcum));
        size_t i = 0; uint32_t at = 0;                                       // at: bucket offset reached so far
        for (uint32_t c = 0; c < n_sub; ++c) {
            const uint32_t lo_c = c << ss, hi_c = (c + 1) << ss;
            uint32_t cc[PGI_RUNS_PER_BLOCK], ce[PGI_RUNS_PER_BLOCK]; unsigned k = 0;
            if (c) { cc[0] = PGI_CODE_UNUSED; ce[0] = lo_c; k = 1; }
            uint64_t add[6] = {0, 0, 0, 0, 0, 0};
            while (i < ne && at < hi_c) {
                const uint32_t e = end[i] < hi_c ? end[i] : hi_c;
                if (k >= PGI_RUNS_PER_BLOCK) throw FormatError("internal: sub-bucket overflow");
                cc[k] = code[i]; ce[k] = e; ++k; add[code[i]] += e - at; at = e;
                if (end[i] <= hi_c) ++i;
            }
            write_block(im.blocks.data() + (size_t) (base + c) * PGI_BLOCK_WORDS, ccum, cc, ce, k, 0u);
            for (unsigned q = 0; q < 6; ++q) ccum[q] += add[q];
        }
        dst = im.blocks.data() + (size_t) b * PGI_BLOCK_WORDS;             // (the vector may have moved)
        write_block(dst, bcum, nullptr, nullptr, 0, (uint32_t) base);    // header: counts before the bucket (the answer at offset 0)
        uint32_t w[PGI_BLOCK_U32];
        for (unsigned j = 0; j < PGI_BLOCK_WORDS; ++j) { w[2 * j] = (uint32_t) dst[j]; w[2 * j + 1] = (uint32_t) (dst[j] >> 32); }
        w[8] &= 0xFFFFu; w[9] = PGI_CODE_DIRBLOCK; w[10] = ss;            // marker; the scan of offset 0 adds nothing whatever the entries
        for (unsigned j = 11; j < PGI_BLOCK_U32; ++j) w[j] = 0;
        for (unsigned j = 0; j < PGI_BLOCK_WORDS; ++j) dst[j] = (uint64_t) w[2 * j] | (uint64_t) w[2 * j + 1] << 32;
        ++n_dir;
    }
    if (pos != n || carry_len || ri != runs.size()) throw FormatError("run lengths do not add up to the BWT size");
    im.k.n_blocks = im.blocks.size() / PGI_BLOCK_WORDS; im.k.n_dir_blocks = n_dir;
    if (im.k.n_blocks >= 0xFFFFFFFFull) throw FormatError("too many run blocks (>= 2^32 - 1) for the 32-bit block indices of the MEM kernel");
    for (unsigned c = 0; c < 6; ++c) totals[c] = cum[c];
}

// Per-byte extension descriptors: SURVEY.md App. B2/B3 folded into weights over true ranks
// (DESIGN.md "Quirk-exact extension"). slot[i] says where R(pos)[i] comes from.
inline void build_descriptors(RiImage &im, const bool present[6], bool legacy) {
    RiConsts &k = im.k;
    const unsigned sigma = (unsigned) k.sigma;
    unsigned slot[8];
    for (unsigned i = 0; i < 8; ++i) slot[i] = PGI_SRC_ZERO;
    k.has_quirk = 0;
    for (unsigned i = 0; i < sigma && i < 6; ++i) {
        if (legacy) {
            // R[i] = cnt_b(nuc[i], rel) + cum_b[sym_map[nuc[i]]]          (r-index.cpp:593-603)
            if (present[i]) slot[i] = i; else { slot[i] = PGI_SRC_QUIRK; k.has_quirk = 1; }
        } else {
            // R[i] = full[sym_map[nuc[i]]], full[c] = true rank of nuc[c] or 0 when absent (r-index.cpp:619-641, 65-82)
            unsigned col = k.sym_map[kNuc[i]];
            slot[i] = (col < 6 && present[col]) ? col : PGI_SRC_ZERO;
        }
    }
    for (unsigned i = 0; i < 8; ++i) k.slot[i] = slot[i];
    uint8_t comp[256];
    for (int i = 0; i < 256; ++i) comp[i] = (uint8_t) i;                   // r-index.cpp:1512-1529
    comp['A'] = 'T'; comp['C'] = 'G'; comp['G'] = 'C'; comp['T'] = 'A';
    comp['a'] = 't'; comp['c'] = 'g'; comp['g'] = 'c'; comp['t'] = 'a';
    uint32_t desc[256];
    for (int a = 0; a < 256; ++a) {
        unsigned w[8] = {0, 0, 0, 0, 0, 0, 0, 0};                          // weights per source
        const unsigned ci = k.sym_map[comp[a]];
        for (unsigned b = 0; b < 6 && k.sym_map[kNuc[b]] < ci; ++b)         // r-index.cpp:733-740 (a while loop: stops at the first failure)
            w[slot[k.sym_map[comp[kNuc[b]]]]]++;
        const unsigned mi = k.sym_map[a];
        uint32_t d = 0;
        for (unsigned c = 0; c < 6; ++c) d |= (uint32_t) w[c] << (4 * c);    // at most 6 increments in total: fits 3 bits
        d |= (uint32_t) w[PGI_SRC_QUIRK] << 24;
        if (slot[mi] < 6) d |= 8u << (4 * slot[mi]);
        else if (slot[mi] == PGI_SRC_QUIRK) d |= PGI_DESC_MAIN_QUIRK;
        else d |= PGI_DESC_MAIN_ZERO;
        d |= (uint32_t) mi << 29;
        desc[a] = d;
    }
    for (int a = 0; a < 256; ++a) { k.desc_b[a] = desc[a]; k.desc_f[a] = desc[comp[a]]; }
    // LF-step descriptors of the whole-read search (pgi_layout.h: desc_c).
    //   legacy  LF (r-index.cpp:650-686): a byte with sym_map 0 fails; otherwise rankAt = true rank of the byte
    //   encoded LF_encoded (:689-711):    rankAt_encoded ranks code symbol_to_code(a) (bytes outside nuc alias code 0,
    //                                     r-index.hpp:664-667) and adds cum[code]: the true rank when six cumulatives exist
    k.count_ok = legacy || sigma == 6;
    for (int a = 0; a < 256; ++a) {
        unsigned code = 0; bool in_nuc = false;
        for (unsigned c = 0; c < 6; ++c) if (kNuc[c] == (uint8_t) a) { code = c; in_nuc = true; }
        const unsigned mi = k.sym_map[a];
        uint32_t d = (uint32_t) mi << 29;
        if (legacy) {
            if (mi != 0 && in_nuc && present[code]) d |= 8u << (4 * code); else d |= PGI_DESC_MAIN_ZERO;
        } else {
            if (present[code]) d |= 8u << (4 * code); else d |= PGI_DESC_MAIN_ZERO;
        }
        k.desc_c[a] = d;
    }
}

// ---------------------------------------------------------------------------------- sparse rank
// Sorted set of positions in [0, size) -> bucket directory + per-element low parts.
// for_each(f) must call f(j, position) for j = 0 .. ones - 1 in increasing order of position.
template <typename ForEach>
inline void build_sparse_rank_from(uint64_t size, uint64_t ones, ForEach &&for_each, SparseRankImage &im) {
    SparseRankConsts &k = im.k; std::memset(&k, 0, sizeof(k));
    k.size = size; k.m = ones;
    unsigned S = 0;                              // about four elements per bucket
    if (ones) while (S < 62 && (size >> S) > ones / 4 + 1) ++S;
    else S = 62;
    k.shift = S; k.low_bytes = S <= 8 ? 1 : S <= 16 ? 2 : S <= 32 ? 4 : 8;
    const uint64_t nbuckets = (size >> S) + 1;
    im.dir.assign(nbuckets + 1, 0);
    im.low.assign((size_t) (ones * k.low_bytes + 16), 0);   // 16 spare bytes: vector loads may over-read
    uint64_t prev = 0; bool first = true;
    for_each([&](uint64_t j, uint64_t p) {
        if (p >= size || (!first && p <= prev)) throw FormatError("sd_vector positions are not strictly increasing inside the universe");
        first = false; prev = p;
        im.dir[(p >> S) + 1]++;
        uint64_t lowv = p & ((S >= 64) ? ~0ULL : ((1ULL << S) - 1));
        std::memcpy(&im.low[(size_t) (j * k.low_bytes)], &lowv, k.low_bytes);
    });
    for (uint64_t b = 0; b < nbuckets; ++b) im.dir[b + 1] += im.dir[b];
    k.n_buckets = nbuckets;
}
inline void build_sparse_rank(const SdVectorView &sd, SparseRankImage &im) {
    build_sparse_rank_from(sd.size, sd.ones, [&](auto &&f) { sd.for_each_one(f); }, im);
}

// ---------------------------------------------------------------------------------- .ri
inline void flatten_ri(const void *bytes, size_t len, RiImage &im) {
    Reader r(bytes, len);
    if (r.get<uint32_t>() != 0x6B3741D8u) throw FormatError("FastLocate: Invalid tag");       // r-index.hpp:85-115
    uint32_t version = r.get<uint32_t>();
    if (version != 1) throw FormatError("FastLocate: Expected v1, got v" + std::to_string(version));
    const uint64_t max_length = r.get<uint64_t>();
    const uint64_t flags = r.get<uint64_t>();
    IntVectorView iv, samples_v, ltr_v; SdVectorView last_v;
    samples_v.parse(r, 0); last_v.parse(r); ltr_v.parse(r, 0);   // not on the find_mems path: kept for locate (below)
    iv.parse(r, 8);
    if (iv.size() != 256) throw FormatError("sym_map must have 256 entries");
    RiConsts &k = im.k; std::memset(&k, 0, sizeof(k));
    for (int i = 0; i < 256; ++i) k.sym_map[i] = (uint8_t) iv.at(i);
    iv.parse(r, 64);
    k.sigma = iv.size();
    if (k.sigma < 1 || k.sigma > 6) throw FormatError("alphabet size outside [1,6] is not supported (nuc has 6 symbols)");
    for (uint64_t i = 0; i < k.sigma; ++i) k.C[i] = iv.at(i);
    for (int i = 0; i < 256; ++i) if (k.sym_map[i] >= k.sigma) throw FormatError("sym_map entry outside the alphabet");
    SdVectorView bsp; bsp.parse(r);
    k.n = r.get<uint64_t>();
    im.n_ref_blocks = bsp.ones; im.encoded = (flags & 1) != 0;
    std::vector<uint64_t> ref_starts; ref_starts.reserve(bsp.ones);
    bsp.for_each_one([&](uint64_t, uint64_t p) { ref_starts.push_back(p); });

    RunCollector bb;
    im.n_file_runs = 0;
    const bool want_locate = samples_v.size() > 0 && max_length > 0;
    std::vector<uint64_t> run_start_pos;            // first row of every LOGICAL run (file record; never merged: each has its own sample)
    if (want_locate) run_start_pos.reserve((size_t) samples_v.size());
    unsigned code_of[256]; for (int i = 0; i < 256; ++i) code_of[i] = 8;
    for (unsigned c = 0; c < 6; ++c) code_of[kNuc[c]] = c;
    auto check_block = [&](uint64_t b, const uint64_t *file_cum, unsigned ncum) {
        if (b >= ref_starts.size() || ref_starts[b] != bb.lpos) throw FormatError("blocks_start_pos disagrees with the run blocks");
        for (unsigned c = 0; c < 6; ++c) {           // file cumulatives are in sym_map order; ours are per nuc code
            if (!bb.lcum[c]) continue;               // an absent symbol has no column of its own
            unsigned col = k.sym_map[kNuc[c]];
            if (col >= ncum || file_cum[col] != bb.lcum[c]) throw FormatError("block cumulative counts disagree with the runs");
        }
    };
    if (!im.encoded) {
        const uint64_t nblk = r.get<uint64_t>();
        for (uint64_t b = 0; b < nblk; ++b) {
            IntVectorView cum; cum.parse(r, 64);
            uint64_t fc[8] = {0}; unsigned nc = (unsigned) (cum.size() < 8 ? cum.size() : 8);
            for (unsigned i = 0; i < nc; ++i) fc[i] = cum.at(i);
            const uint64_t nr = r.get<uint64_t>();
            if (nr) { check_block(b, fc, nc); im.q_val.push_back(fc[0]); }
            for (uint64_t j = 0; j < nr; ++j) {
                uint64_t sym = r.get<uint64_t>(), l = r.get<uint64_t>();
                if (sym > 255 || code_of[sym] > 5) throw FormatError("BWT symbol outside {\\n,A,C,G,N,T} is not supported");
                if (want_locate && l) run_start_pos.push_back(bb.lpos);
                bb.add(code_of[sym], l); ++im.n_file_runs;
            }
        }
        im.q_start = ref_starts;
    } else {
        const uint64_t ebs = r.get<uint64_t>(); (void) ebs;
        const bool hasN = r.get<uint8_t>() != 0;
        const unsigned cum_len = hasN ? 6 : 5;                                  // r-index.cpp:630
        if (k.sigma != cum_len) throw FormatError("encoded r-index whose alphabet size disagrees with its N flag (unreadable by the reference too)");
        IntVectorView offs; offs.parse(r, 0);
        const uint64_t nbytes = r.get<uint64_t>();
        const uint8_t *stream = r.take(nbytes);
        const uint64_t nblk = bsp.ones;                                         // blocks without a start bit are never reached
        if (offs.size() < nblk) throw FormatError("encoded block offsets are shorter than blocks_start_pos");
        for (uint64_t b = 0; b < nblk; ++b) {
            const uint8_t *p = stream + offs.at(b);
            const uint8_t *end = (b + 1 < offs.size()) ? stream + offs.at(b + 1) : stream + nbytes;
            if (p > end || end > stream + nbytes) throw FormatError("encoded block offsets out of range");
            uint64_t fc[8] = {0};
            for (unsigned i = 0; i < cum_len; ++i) fc[i] = bytecode_next(p, end);
            check_block(b, fc, cum_len);
            while (p < end) {
                uint8_t h = *p++; unsigned code = h >> 5; uint64_t l = (h & 31) + 1;
                if ((h & 31) == 31) l = 32 + bytecode_next(p, end);
                if (code > 5) throw FormatError("encoded run with symbol code > 5");
                if (want_locate) run_start_pos.push_back(bb.lpos);
                bb.add(code, l); ++im.n_file_runs;
            }
        }
    }
    bb.finish();
    if (bb.lpos != k.n) throw FormatError("run lengths do not add up to the BWT size");
    uint64_t totals[6];
    build_buckets(im, bb.runs, k.n, totals);
    bool present[6]; for (unsigned c = 0; c < 6; ++c) present[c] = totals[c] != 0;
    build_descriptors(im, present, !im.encoded);
    // locate: samples[run] = SA value at the first row of logical run `run`; phi through last / last_to_run
    // (src/r-index.cpp:1300-1343, 1369-1372). On an encoded file without N the reference's run lookup skips six
    // ByteCode values where five exist (:85-90) -- refused like count (pgi_layout.h).
    k.max_length = max_length; k.locate_ok = 0;
    if (want_locate && run_start_pos.size() == samples_v.size() && last_v.ones == ltr_v.size() && (!im.encoded || k.sigma == 6)) {
        build_sparse_rank_from(k.n, run_start_pos.size(), [&](auto &&f) { for (uint64_t j = 0; j < run_start_pos.size(); ++j) f(j, run_start_pos[j]); }, im.run_starts);
        build_sparse_rank(last_v, im.last);
        pack_from_view(samples_v, im.samples); pack_from_view(ltr_v, im.last_to_run);
        k.locate_ok = 1;
    }
    if (!k.has_quirk) { im.q_start.clear(); im.q_val.clear(); }
    k.n_quirk = im.q_start.size();
    k.mem_tail = r.remaining();   // bytes after the index (0 for a well-formed file)
}

// ---------------------------------------------------------------------------------- tags
inline void flatten_tags(const void *bytes, size_t len, int layout, TagImage &im) {
    Reader r(bytes, len);
    im.layout = layout;
    if (layout == 0) {
        IntVectorView items; items.parse(r, 0);
        if (items.width == 0 || items.width > 64) throw FormatError("tag item width outside [1,64]");
        im.n_items = items.size(); im.width = items.width;
        size_t nw = (size_t) ((items.bits + 63) / 64);
        im.words.assign(nw + 2, 0);
        std::memcpy(im.words.data(), items.payload, nw * 8);
    } else if (layout == 1) {
        uint64_t nbytes = r.get<uint64_t>();
        const uint8_t *p = r.take(nbytes);
        im.stream.assign(p, p + nbytes);
        im.n_items = nbytes; im.width = 8;
        im.stream.resize((size_t) nbytes + 16, 0);      // reads past the end see zero bytes (defined UB)
    } else throw FormatError("unknown tag layout");
    SdVectorView starts; starts.parse(r);
    im.starts.reserve(starts.ones);
    im.starts_regular = layout == 0;
    starts.for_each_one([&](uint64_t j, uint64_t p) { im.starts.push_back(p); if (p != 10 * j) im.starts_regular = false; });
    SdVectorView iv; iv.parse(r);
    if (r.remaining()) throw FormatError("trailing bytes after the tag arrays (wrong tag layout?)");
    build_sparse_rank(iv, im.intervals);
}

} // namespace pgi
