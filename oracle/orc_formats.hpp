// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement of the on-disk layouts the reference's find_mems path reads.
// Nothing under oracle/ is linked, imported or executed by the product path
// (pangenome-index_b200/); only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may use it, and only as the checker.
//
// Parity status: the reference cannot be compiled here (SDSL-vgteam, GBWT,
// GBWTGraph, libhandlegraph are neither vendored nor installed), so these
// readers are pinned by consuming every shipped binary fixture to its last
// byte and reproducing the values recorded in SURVEY.md App. D
// (tests/test_oracle_formats.py).
//
// Third-party layouts restated (un-vendored, unpinned in the reference):
//   sdsl::int_vector<0>/<8>/<64>, sdsl::bit_vector, sdsl::sd_vector<> (+ the two
//   serialized select_support_mcl blobs, skipped structurally), gbwt::ByteCode.
// Reference call sites: src/r-index.cpp:378-459 (FastLocate::load/load_encoded),
//   include/pangenome_index/r-index.hpp:283-292 (Run_blocks::load),
//   src/tag_arrays.cpp:735-776 (TagArray loaders).
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include <algorithm>

namespace orc {

struct Cursor {
    const uint8_t *p; size_t len; size_t off;
    Cursor(const uint8_t *p_, size_t n) : p(p_), len(n), off(0) {}
    void need(size_t n) const {
        if (off + n > len || off + n < off) throw std::runtime_error("oracle: truncated input at byte " + std::to_string(off));
    }
    uint64_t u64() { need(8); uint64_t v; std::memcpy(&v, p + off, 8); off += 8; return v; }
    uint32_t u32() { need(4); uint32_t v; std::memcpy(&v, p + off, 4); off += 4; return v; }
    uint8_t u8() { need(1); return p[off++]; }
    const uint8_t *bytes(size_t n) { need(n); const uint8_t *r = p + off; off += n; return r; }
    bool eof() const { return off == len; }
};

// sdsl::int_vector<w>: u64 size_in_bits, [u8 width iff w==0], ceil(bits/64) u64 words.
struct IntVec {
    uint64_t bits = 0; uint8_t width = 64; std::vector<uint64_t> words;
    uint64_t size() const { return width ? bits / width : 0; }
    bool empty() const { return size() == 0; }
    uint64_t get(uint64_t i) const {
        uint64_t b = i * width, w = b >> 6, s = b & 63;
        uint64_t v = words[w] >> s;
        if (s + width > 64) v |= words[w + 1] << (64 - s);
        return width == 64 ? v : (v & ((1ULL << width) - 1));
    }
    void load(Cursor &c, unsigned fixed_width) {
        bits = c.u64();
        width = fixed_width ? (uint8_t) fixed_width : c.u8();
        size_t nw = (size_t) ((bits + 63) >> 6);
        const uint8_t *src = c.bytes(nw * 8);
        words.assign(nw + 1, 0);               // one spare word keeps get() branch-free at the tail
        std::memcpy(words.data(), src, nw * 8);
    }
};

// sdsl::select_support_mcl<b,1> serialized blob: skipped, never interpreted.
inline void skip_select_support(Cursor &c) {
    uint64_t arg_cnt = c.u64();
    if (arg_cnt == 0) return;
    IntVec tmp; tmp.load(c, 0);                 // superblock
    tmp.load(c, 1);                             // mini_or_long
    uint64_t sb = (arg_cnt + 4095) >> 12;
    for (uint64_t i = 0; i < sb; ++i) tmp.load(c, 0);   // one long or mini vector per superblock
}

// sdsl::sd_vector<>: materialised as the sorted list of 1-positions plus a bucket
// directory (array-based rank/select/predecessor; results identical to SDSL's).
struct SdVec {
    uint64_t size = 0; std::vector<uint64_t> ones;
    unsigned dshift = 0; std::vector<uint64_t> dir;    // dir[b] = #ones with position < (b << dshift)

    void load(Cursor &c) {
        size = c.u64();
        uint8_t wl = c.u8();
        IntVec low, high;
        low.load(c, 0);
        high.load(c, 1);
        skip_select_support(c);
        skip_select_support(c);
        uint64_t m = (low.width == 0) ? 0 : low.size();
        // wl==0 is legal (low has width 0 and no payload): count ones from high instead.
        ones.clear();
        uint64_t j = 0;
        for (uint64_t w = 0; w * 64 < high.bits; ++w) {
            uint64_t x = high.words[w];
            while (x) {
                unsigned b = __builtin_ctzll(x); x &= x - 1;
                uint64_t pbit = w * 64 + b;
                if (pbit >= high.bits) break;
                uint64_t lo = (wl == 0) ? 0 : low.get(j);
                ones.push_back(((pbit - j) << wl) | lo);
                ++j;
            }
        }
        if (wl != 0 && j != m) throw std::runtime_error("oracle: sd_vector low/high mismatch");
        build_dir();
    }
    void build_dir() {
        uint64_t m = ones.size();
        dshift = 0;
        if (m == 0) { dir.assign(2, 0); dshift = 63; return; }
        // about 2 ones per bucket
        while (dshift < 62 && ((size >> dshift) > m / 2 + 1)) ++dshift;
        uint64_t nb = (size >> dshift) + 2;
        dir.assign(nb + 1, 0);
        uint64_t j = 0;
        for (uint64_t b = 0; b <= nb; ++b) {
            uint64_t lim = b << dshift;
            while (j < m && ones[j] < lim) ++j;
            dir[b] = j;
        }
    }
    // #ones in [0, i)
    uint64_t rank1(uint64_t i) const {
        if (i >= size) return i > size ? ones.size() : (uint64_t) (std::lower_bound(ones.begin(), ones.end(), i) - ones.begin());
        uint64_t b = i >> dshift;
        uint64_t lo = dir[b], hi = dir[b + 1];
        return (uint64_t) (std::lower_bound(ones.begin() + lo, ones.begin() + hi, i) - ones.begin());
    }
    // i is 1-based; out of range -> UINT64_MAX (undefined in the reference)
    uint64_t select1(uint64_t i) const { return (i >= 1 && i <= ones.size()) ? ones[i - 1] : UINT64_MAX; }
    // last one at position <= i : returns index j (or UINT64_MAX if none)
    uint64_t predecessor_index(uint64_t i) const {
        uint64_t r = (i >= size) ? ones.size() : rank1(i + 1);
        return r == 0 ? UINT64_MAX : r - 1;
    }
};

// gbwt::ByteCode::read : 7 data bits per byte, low group first, bit 7 = continue.
// Reading past the end of the array is undefined in the reference; here every
// byte past the end reads as 0x00 (documented extension, DESIGN.md "defined UB").
inline uint64_t bytecode_read(const uint8_t *a, size_t n, size_t &i) {
    uint64_t res = 0; unsigned shift = 0;
    for (;;) {
        uint8_t b = (i < n) ? a[i] : 0; ++i;
        if (shift < 64) res |= (uint64_t) (b & 0x7F) << shift;
        shift += 7;
        if (!(b & 0x80)) break;
    }
    return res;
}
inline void bytecode_write(std::vector<uint8_t> &out, uint64_t v) {
    while (v > 0x7F) { out.push_back((uint8_t) ((v & 0x7F) | 0x80)); v >>= 7; }
    out.push_back((uint8_t) v);
}

} // namespace orc
