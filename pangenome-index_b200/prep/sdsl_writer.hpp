// Byte-exact writers for the SDSL (vgteam fork) on-disk layouts the reference's .ri / .tags
// files are made of, so that indexes built here load in the real reference:
//   int_vector<0>/<8>/<64>, bit_vector, sd_vector<> including its two select_support_mcl blobs.
// Layouts follow SURVEY.md App. A0 and were checked by regenerating the shipped
// test_data/bidirectional_test/xy.ri byte for byte from contigs_xy.rl_bwt (tests/test_prep.py).
// Reference writers being replaced: FastLocate::serialize / serialize_encoded
// (src/r-index.cpp:266-376), TagArray::merge_compressed_files_sdsl (src/tag_arrays.cpp:622-654).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

namespace pgi_prep {

struct ByteSink {
    std::vector<uint8_t> buf;
    void u64(uint64_t v) { size_t o = buf.size(); buf.resize(o + 8); std::memcpy(&buf[o], &v, 8); }
    void u32(uint32_t v) { size_t o = buf.size(); buf.resize(o + 4); std::memcpy(&buf[o], &v, 4); }
    void u8(uint8_t v) { buf.push_back(v); }
    void bytes(const void *p, size_t n) { size_t o = buf.size(); buf.resize(o + n); if (n) std::memcpy(&buf[o], p, n); }
};

static inline unsigned bits_hi(uint64_t x) { return x ? 63u - (unsigned) __builtin_clzll(x) : 0u; }   // sdsl::bits::hi
static inline unsigned bits_length(uint64_t x) { return x ? bits_hi(x) + 1 : 0; }                      // sdsl::bits::length

// Packed little-endian bit array with fixed element width (sdsl::int_vector<0> payload).
struct PackedVec {
    uint64_t n = 0; unsigned width = 1; std::vector<uint64_t> words;
    PackedVec() {}
    PackedVec(uint64_t n_, unsigned w) { init(n_, w); }
    void init(uint64_t n_, unsigned w) { n = n_; width = w; words.assign((size_t) ((n * w + 63) >> 6) + 1, 0); }
    void set(uint64_t i, uint64_t v) {
        if (width == 0) return;
        uint64_t b = i * width, w = b >> 6, s = b & 63;
        uint64_t mask = width == 64 ? ~0ULL : ((1ULL << width) - 1);
        v &= mask;
        words[w] = (words[w] & ~(mask << s)) | (v << s);
        if (s + width > 64) { unsigned r = 64 - (unsigned) s; words[w + 1] = (words[w + 1] & ~(mask >> r)) | (v >> r); }
    }
    uint64_t get(uint64_t i) const {
        if (width == 0) return 0;
        uint64_t b = i * width, w = b >> 6, s = b & 63;
        uint64_t v = words[w] >> s;
        if (s + width > 64) v |= words[w + 1] << (64 - s);
        return width == 64 ? v : (v & ((1ULL << width) - 1));
    }
    // int_vector<0>: u64 bits, u8 width, words
    void serialize_iv0(ByteSink &out) const {
        uint64_t bits = n * width; out.u64(bits); out.u8((uint8_t) width);
        out.bytes(words.data(), (size_t) ((bits + 63) >> 6) * 8);
    }
    // int_vector<W> with compile-time width: no width byte
    void serialize_fixed(ByteSink &out) const {
        uint64_t bits = n * width; out.u64(bits);
        out.bytes(words.data(), (size_t) ((bits + 63) >> 6) * 8);
    }
};

// select_support_mcl<b,1>::serialize over bit_vector `bv` of `nbits` bits (b = which bit value is selected).
static inline void write_select_support(ByteSink &out, const std::vector<uint64_t> &bv, uint64_t nbits, int b) {
    const uint64_t SB = 4096;
    uint64_t arg_cnt = 0;
    for (uint64_t w = 0; w * 64 < nbits; ++w) {
        uint64_t x = b ? bv[w] : ~bv[w];
        uint64_t rem = nbits - w * 64;
        if (rem < 64) x &= (1ULL << rem) - 1;
        arg_cnt += (uint64_t) __builtin_popcountll(x);
    }
    out.u64(arg_cnt);
    if (arg_cnt == 0) return;
    uint64_t capacity = ((nbits + 63) >> 6) << 6;
    unsigned logn = bits_hi(capacity) + 1;
    uint64_t logn4 = (uint64_t) logn * logn * logn * logn;
    uint64_t sb = (arg_cnt + SB - 1) / SB;
    PackedVec superblock(sb, logn);
    std::vector<PackedVec> blocks(sb);
    std::vector<uint8_t> is_mini(sb, 0);
    bool any_long = false;
    std::vector<uint64_t> argpos(SB);
    uint64_t cnt = 0, sbi = 0;
    for (uint64_t w = 0; w * 64 < nbits; ++w) {
        uint64_t x = b ? bv[w] : ~bv[w];
        uint64_t rem = nbits - w * 64;
        if (rem < 64) x &= (1ULL << rem) - 1;
        while (x) {
            unsigned bit = (unsigned) __builtin_ctzll(x); x &= x - 1;
            argpos[cnt % SB] = w * 64 + bit;
            ++cnt;
            if (cnt % SB == 0 || cnt == arg_cnt) {
                uint64_t lasti = (cnt - 1) % SB;
                superblock.set(sbi, argpos[0]);
                uint64_t pos_diff = argpos[lasti] - argpos[0];
                if (pos_diff > logn4) {
                    any_long = true;
                    blocks[sbi].init(SB, bits_hi(argpos[lasti]) + 1);
                    for (uint64_t j = 0; j <= lasti; ++j) blocks[sbi].set(j, argpos[j]);
                } else {
                    is_mini[sbi] = 1;
                    blocks[sbi].init(64, bits_hi(pos_diff) + 1);
                    for (uint64_t j = 0; j <= lasti; j += 64) blocks[sbi].set(j / 64, argpos[j] - argpos[0]);
                }
                ++sbi;
            }
        }
    }
    superblock.serialize_iv0(out);
    // mini_or_long: empty bit_vector unless at least one long superblock exists
    if (any_long) {
        PackedVec mol(sb, 1);
        for (uint64_t i = 0; i < sb; ++i) mol.set(i, is_mini[i]);
        mol.serialize_fixed(out);
    } else {
        out.u64(0);
    }
    for (uint64_t i = 0; i < sb; ++i) blocks[i].serialize_iv0(out);
}

// sd_vector<> built like sdsl::sd_vector_builder(n, m) followed by set() of sorted positions.
// Parameters follow the simple-sds compatible get_params(): low width = round(max(log2(n ln2 / m), 1)),
// high has m + ceil(n / 2^wl) bits.
static inline void write_sd_vector(ByteSink &out, uint64_t n, const std::vector<uint64_t> &ones) {
    uint64_t m = ones.size();
    unsigned wl = 1;
    if (m > 0 && m <= n) {
        double ideal = std::log2(((double) n * std::log(2.0)) / (double) m);
        wl = (unsigned) std::llround(std::max(ideal, 1.0));
    }
    uint64_t buckets = (n >> wl) + ((n & ((1ULL << wl) - 1)) ? 1 : 0);
    PackedVec low(m, wl);
    uint64_t hbits = m + buckets;
    std::vector<uint64_t> high((size_t) ((hbits + 63) >> 6) + 1, 0);
    for (uint64_t j = 0; j < m; ++j) {
        low.set(j, ones[j] & ((1ULL << wl) - 1));
        uint64_t hp = (ones[j] >> wl) + j;
        high[hp >> 6] |= 1ULL << (hp & 63);
    }
    out.u64(n); out.u8((uint8_t) wl);
    low.serialize_iv0(out);
    out.u64(hbits); out.bytes(high.data(), (size_t) ((hbits + 63) >> 6) * 8);
    write_select_support(out, high, hbits, 1);
    write_select_support(out, high, hbits, 0);
}

static inline void bytecode_write(std::vector<uint8_t> &out, uint64_t v) {
    while (v > 0x7F) { out.push_back((uint8_t) ((v & 0x7F) | 0x80)); v >>= 7; }
    out.push_back((uint8_t) v);
}

} // namespace pgi_prep
