// pgi_core.cuh — the query arithmetic of the find_mems hot path over the flattened device layout.
// Every function here runs on the GPU inside the kernels of pgi_kernels.cu. The header also compiles
// as plain C++ (PGI_HD empty) for tests/hostsim, a test-only single-step harness that lets the lane
// state machine be checked against the oracle on the GPU-less build box; the product library never
// executes it on the CPU.
//
// Reference semantics restated here (file:line relative to the reference tree):
//   rank vector        FastLocate::rank_at_cached[_encoded]        src/r-index.cpp:593-641
//   FMD extension      FastLocate::backward_extend[_encoded]       src/r-index.cpp:713-756, 1395-1428
//                      FastLocate::forward_extend[_encoded]        src/r-index.cpp:758-764, 1500-1509
//   MEM finder         find_mems_function / find_all_mems          include/pangenome_index/algorithm.hpp:653-757
//   tag lookup         TagArray::query_compressed[_compact]        src/tag_arrays.cpp:780-890
#pragma once
#include <stdint.h>
#include "pgi_layout.h"

#ifdef __CUDACC__
#define PGI_HD __host__ __device__ __forceinline__
#else
#define PGI_HD inline
#endif

namespace pgi {

// Device view of the flattened r-index.
struct RiDev {
    const uint64_t *blocks;     // 8 words per 64-byte block: block b = bucket b = positions [b << bshift, (b + 1) << bshift), then chunks
    const uint64_t *q_start;    // quirk tables (n_quirk entries) or nullptr
    const uint64_t *q_val;
    uint64_t n, n_blocks, n_quirk;
    uint32_t bshift, has_quirk;
    uint64_t C[8];
    uint64_t n_buckets_dbg;     // primary (bucket) blocks; chunk blocks follow (bounds-checking build only)
};

struct SparseRankDev {
    const uint64_t *dir; const uint8_t *low;
    uint64_t size, m; uint32_t shift, low_bytes;
};

// Bit-packed int_vector<0> as in the file; locate structures (toehold samples + phi).
struct PackedDev { const uint64_t *words; uint64_t n; uint32_t width, pad; };
struct LocDev {
    SparseRankDev run_starts, last;     // first BWT row of every logical run; FastLocate::last
    PackedDev samples, last_to_run;
    uint64_t max_length;
};

struct TagDev {
    SparseRankDev iv;           // bwt_intervals
    const uint64_t *starts; uint64_t n_starts;
    const uint64_t *words;      // sdsl-compact items
    const uint8_t *stream;      // ByteCode stream
    uint64_t n_items;           // items (layout 0) or stream bytes (layout 1)
    uint32_t width, layout, starts_regular, pad;
};

// ------------------------------------------------------------------------------------ loads
PGI_HD uint64_t ld64(const uint64_t *p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}
// One 64-byte run block as sixteen u32 words (four 128-bit loads of one aligned 64 B line).
PGI_HD void ld_block(const uint64_t *blk, uint32_t w[16]) {
#if defined(__CUDA_ARCH__)
    const uint4 *q = reinterpret_cast<const uint4 *>(blk);
    const uint4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    w[8] = c.x; w[9] = c.y; w[10] = c.z; w[11] = c.w; w[12] = d.x; w[13] = d.y; w[14] = d.z; w[15] = d.w;
#else
    for (int i = 0; i < 8; ++i) { w[2 * i] = (uint32_t) blk[i]; w[2 * i + 1] = (uint32_t) (blk[i] >> 32); }
#endif
}
PGI_HD uint32_t ld32(const uint32_t *p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}

// ------------------------------------------------------------------------------------ in-block rank
// ---- SIMD primitives of the scan (one SASS instruction each on sm_100a; plain C++ for the test-only host build)
PGI_HD uint32_t min_u16x2(uint32_t a, uint32_t b) {                 // VIMNMX.U16x2
#if defined(__CUDA_ARCH__)
    return __vminu2(a, b);
#else
    const uint32_t al = a & 0xFFFFu, bl = b & 0xFFFFu, ah = a >> 16, bh = b >> 16;
    return (al < bl ? al : bl) | (ah < bh ? ah : bh) << 16;
#endif
}
// Selector nibbles are run codes: 0..7 in every block that is scanned with a non-zero offset (pgi_layout.h), so the raw PRMT
// is used (__byte_perm masks the selector with 0x7777 first: 14 LOP3 per extension). The 0xF marker nibble of a directory
// block selects a sign-replicated byte, and that block is only ever scanned at offset 0, where every product is 0 * byte.
PGI_HD uint32_t byte_perm(uint32_t lo, uint32_t hi, uint32_t sel) { // PRMT: result byte i = byte (sel nibble i) of {hi:lo}
#if defined(__CUDA_ARCH__)
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(hi), "r"(sel));
    return r;
#else
    const uint64_t t = (uint64_t) hi << 32 | lo; uint32_t r = 0;
    for (int i = 0; i < 4; ++i) r |= (uint32_t) ((t >> (8 * ((sel >> (4 * i)) & 7u))) & 0xFFu) << (8 * i);
    return r;
#endif
}
PGI_HD uint32_t dp2a_lo(uint32_t a, uint32_t b, uint32_t c) {       // IDP.2A.LO.U16.U8: a.lo16 * b.byte0 + a.hi16 * b.byte1 + c
#if defined(__CUDA_ARCH__)
    return __dp2a_lo(a, b, c);
#else
    return (a & 0xFFFFu) * (b & 0xFFu) + (a >> 16) * ((b >> 8) & 0xFFu) + c;
#endif
}
PGI_HD uint32_t dp2a_hi(uint32_t a, uint32_t b, uint32_t c) {       // IDP.2A.HI.U16.U8: a.lo16 * b.byte2 + a.hi16 * b.byte3 + c
#if defined(__CUDA_ARCH__)
    return __dp2a_hi(a, b, c);
#else
    return (a & 0xFFFFu) * ((b >> 16) & 0xFFu) + (a >> 16) * (b >> 24) + c;
#endif
}

PGI_HD uint32_t dp4a_u8(uint32_t a, uint32_t b, uint32_t c) {       // IDP.4A.U8.U8: sum of the four byte products + c
#if defined(__CUDA_ARCH__)
    return __dp4a(a, b, c);
#else
    for (int i = 0; i < 4; ++i) c += ((a >> (8 * i)) & 0xFFu) * ((b >> (8 * i)) & 0xFFu);
    return c;
#endif
}

// Byte tables of a descriptor's six 4-bit fields (bits 0..2 weight w[c], bit 3 "main symbol"), indexed by code through PRMT:
// T.x / T.y = weights of codes 0..3 / 4..7, T.z / T.w = main flags of codes 0..3 / 4..7 (codes 6, 7: zero).
struct ScanTab { uint32_t x, y, z, w; };
PGI_HD ScanTab scan_tab(uint32_t tbl) {
    ScanTab t;
    t.x = (tbl & 7u) | ((tbl >> 4) & 7u) << 8 | ((tbl >> 8) & 7u) << 16 | ((tbl >> 12) & 7u) << 24;
    t.y = ((tbl >> 16) & 7u) | ((tbl >> 20) & 7u) << 8;
    t.z = ((tbl >> 3) & 1u) | ((tbl >> 7) & 1u) << 8 | ((tbl >> 11) & 1u) << 16 | ((tbl >> 15) & 1u) << 24;
    t.w = ((tbl >> 19) & 1u) | ((tbl >> 23) & 1u) << 8;
    return t;
}

// Weighted rank of one position. tbl = the low 24 bits of an extension descriptor, T = scan_tab(tbl). Returns, for the
// first `rel` positions of the block plus the block's cumulative header:
//     W = sum_c w[c] * rank_c(pos)      M = rank_main(pos)   (0 when no field has bit 3).
// In-block part: with m_i = min(rel, E_i) (pgi_layout.h), run i contributes (m_i - m_{i-1}) * weight(c_i), so
//     sum_i m_i * weight(c_i)  -  sum_i m_i * weight(c_{i+1})          (weight(c_12) = 0)
// : per four runs two 16x2 mins, four PRMT table lookups (codes and next codes as selectors; weights and main flags) and
// eight dp2a. No dependent chain over the runs, no per-run shifts and masks.
struct PosRank { uint64_t W, M; };

// rel = offset of the position inside its bucket (< 2^bshift <= 256)
PGI_HD PosRank pos_rank(const uint32_t w[16], uint32_t rel, uint32_t tbl, const ScanTab &T) {
    const uint32_t rr = rel * 0x10001u;
    const uint32_t c0 = w[9], c1 = w[8] >> 16;                          // codes 0..7, 8..11
    const uint32_t n0 = c0 >> 4 | c1 << 28, n1 = c1 >> 4 | 0x7000u;     // next codes; behind the last run: unused (weight 0)
    uint32_t aw = 0, an = 0, fw = 0, fn = 0, m1 = 0;
#pragma unroll
    for (int g = 0; g < 3; ++g) {
        const uint32_t m0 = min_u16x2(rr, w[10 + 2 * g]);
        m1 = min_u16x2(rr, w[11 + 2 * g]);
        const uint32_t sc = g == 0 ? c0 : g == 1 ? c0 >> 16 : c1, sn = g == 0 ? n0 : g == 1 ? n0 >> 16 : n1;
        const uint32_t bw = byte_perm(T.x, T.y, sc), bn = byte_perm(T.x, T.y, sn);
        const uint32_t bf = byte_perm(T.z, T.w, sc), bg = byte_perm(T.z, T.w, sn);
        aw = dp2a_hi(m1, bw, dp2a_lo(m0, bw, aw)); an = dp2a_hi(m1, bn, dp2a_lo(m0, bn, an));
        fw = dp2a_hi(m1, bf, dp2a_lo(m0, bf, fw)); fn = dp2a_hi(m1, bg, dp2a_lo(m0, bg, fn));
    }
    // header: the six counts before the block in limb form (pgi_layout.h), code order = byte order of the tables T, so the
    // weighted sums are dot products as well: bits 0..15 and 16..31 of the counts three dp2a each, bits 32..39 two dp4a
    // (the code nibbles in the upper half of w8 meet the zero bytes 2, 3 of T.y / T.w). No weight extraction, no 64-bit
    // multiplies: 20 instructions per position for W and M together (the per-count form took 37 plus 12 shared).
    const uint32_t w0 = dp2a_lo(w[3], T.y, dp2a_hi(w[2], T.x, dp2a_lo(w[1], T.x, 0u)));
    const uint32_t w1 = dp2a_lo(w[6], T.y, dp2a_hi(w[5], T.x, dp2a_lo(w[4], T.x, 0u)));
    const uint32_t w2 = dp4a_u8(w[8], T.y, dp4a_u8(w[7], T.x, 0u));
    const uint32_t m0 = dp2a_lo(w[3], T.w, dp2a_hi(w[2], T.z, dp2a_lo(w[1], T.z, 0u)));
    const uint32_t m1h = dp2a_lo(w[6], T.w, dp2a_hi(w[5], T.z, dp2a_lo(w[4], T.z, 0u)));
    const uint32_t m2 = dp4a_u8(w[8], T.w, dp4a_u8(w[7], T.z, 0u));
    (void) tbl;
    PosRank o;
    o.W = (uint64_t) w0 + ((uint64_t) w1 << 16) + ((uint64_t) w2 << 32) + (uint64_t) (aw - an);
    o.M = (uint64_t) m0 + ((uint64_t) m1h << 16) + ((uint64_t) m2 << 32) + (uint64_t) (fw - fn);
    return o;
}
PGI_HD bool is_dir_block(const uint32_t w[16]) { return (w[9] & 0xFu) == PGI_CODE_DIRBLOCK; }

// Crowded bucket (pgi_layout.h): w is its directory block; pick the chunk containing rel and rank there. One more
// dependent 64-byte access; rare by the choice of bshift.
#if defined(__CUDA_ARCH__)
__device__ __noinline__
#else
inline
#endif
PosRank rank_in_chunk(const RiDev &ix, const uint32_t w[16], uint32_t rel, uint32_t tbl) {
    uint32_t v[16];
    ld_block(ix.blocks + ((uint64_t) w[0] + (rel >> w[10])) * PGI_BLOCK_WORDS, v);      // sub-bucket number = rel >> sub_shift
    return pos_rank(v, rel, tbl, scan_tab(tbl));
}
// Rank of one position from memory (batched rank API, lane-private fetch mode, hostsim).
PGI_HD PosRank rank_pos(const RiDev &ix, uint64_t pos, uint32_t tbl) {
    uint32_t w[16];
    ld_block(ix.blocks + (pos >> ix.bshift) * PGI_BLOCK_WORDS, w);
    const uint32_t rel = (uint32_t) pos & ((1u << ix.bshift) - 1u);
    if (rel && is_dir_block(w)) return rank_in_chunk(ix, w, rel, tbl);
    return pos_rank(w, rel, tbl, scan_tab(tbl));     // offset 0 of a crowded bucket: the directory block's header is the answer
}
// True ranks of the six nuc codes at pos (batched rank API and tests; not on the MEM path).
PGI_HD void rank_all(const RiDev &ix, uint64_t pos, uint64_t tr[6]) {
#pragma unroll 1
    for (int c = 0; c < 6; ++c) tr[c] = rank_pos(ix, pos, 8u << (4 * c)).M;
}

// Quirk slot value Q(pos): column-0 cumulative of the reference's own 10-run block containing pos
// (legacy files with an absent nuc symbol only). Binary search over q_start.
PGI_HD uint64_t quirk_at(const RiDev &ix, uint64_t pos) {
    uint64_t lo = 0, up = ix.n_quirk;                // last j with q_start[j] <= pos; q_start[0] == 0
    if (up == 0) return 0;
    --up;
    while (lo < up) {
        uint64_t mid = (lo + up + 1) >> 1;
        if (ld64(ix.q_start + mid) <= pos) lo = mid; else up = mid - 1;
    }
    return ld64(ix.q_val + lo);
}

// ------------------------------------------------------------------------------------ FMD extension
struct BiInt { uint64_t k, kp; int64_t s; };      // FastLocate::bi_interval {forward, reverse, size}

// Second half of an extension: from the two weighted ranks to the new interval (src/r-index.cpp:733-755).
PGI_HD BiInt extend_finish(const RiDev &ix, const BiInt &in, uint32_t d, const PosRank &r1, const PosRank &r2, uint64_t p1, uint64_t p2) {
    uint64_t kp = in.kp + (r2.W - r1.W), x = r2.M, y = r1.M;
    if (ix.has_quirk) {
        const uint32_t wq = (d >> 24) & 7u;
        if (wq || (d & PGI_DESC_MAIN_QUIRK)) {
            const uint64_t q2 = quirk_at(ix, p2), q1 = quirk_at(ix, p1);
            kp += (uint64_t) wq * (q2 - q1);
            if (d & PGI_DESC_MAIN_QUIRK) { x = q2; y = q1; }
        }
    }
    BiInt out;
    if (y >= x) { out.k = 0; out.kp = 0; out.s = 0; return out; }
    out.k = y + ix.C[d >> 29]; out.kp = kp; out.s = (int64_t) (x - y);
    return out;
}

// One backward extension with the per-byte descriptor d (desc_b[a]); forward extension = the same on
// the swapped interval with desc_f[a] (src/r-index.cpp:758-764). All arithmetic mod 2^64, like the
// reference's size_t. Positions beyond n read as n (defined UB).
// Both positions (k and k+s) go through the same straight-line code, whether or not they share a bucket.
PGI_HD BiInt extend_desc(const RiDev &ix, BiInt in, uint32_t d) {
    uint64_t p2 = in.k + (uint64_t) in.s, p1 = in.k;
    if (p2 > ix.n) p2 = ix.n;
    if (p1 > ix.n) p1 = ix.n;
    const uint32_t tbl = d & 0xFFFFFFu, mask = (1u << ix.bshift) - 1u;
    const uint64_t b1 = p1 >> ix.bshift, b2 = p2 >> ix.bshift;
    const ScanTab T = scan_tab(tbl);
    uint32_t w[16];
    ld_block(ix.blocks + b1 * PGI_BLOCK_WORDS, w);
    PosRank r1 = (((uint32_t) p1 & mask) && is_dir_block(w)) ? rank_in_chunk(ix, w, (uint32_t) p1 & mask, tbl) : pos_rank(w, (uint32_t) p1 & mask, tbl, T);
    if (b2 != b1) ld_block(ix.blocks + b2 * PGI_BLOCK_WORDS, w);        // k and k+s in one bucket (the common case once the interval is short): no second load
    PosRank r2 = (((uint32_t) p2 & mask) && is_dir_block(w)) ? rank_in_chunk(ix, w, (uint32_t) p2 & mask, tbl) : pos_rank(w, (uint32_t) p2 & mask, tbl, T);
    return extend_finish(ix, in, d, r1, r2, p1, p2);
}

// ------------------------------------------------------------------------------------ MEM state machine
// One lane walks one read through find_all_mems as a loop whose every trip performs exactly one FMD
// extension (so the 32 lanes of a warp stay convergent although they are in different phases).
enum { PH_DONE = 0, PH_P1 = 1, PH_P2 = 2, PH_P3 = 3, PH_EMIT = 4 };
enum { LANE_DONE = 0, LANE_EXTEND = 1, LANE_EMIT = 2 };

// Read positions are 32-bit: a single read is limited to 2^31 - 16 bytes (longer reads are truncated).
#define PGI_MAX_READ_LEN 0x7FFFFFF0ull
struct MemLane {
    uint32_t len, x, j, phase;
    BiInt b;                    // interval being extended (bint in P1/P2, back in P3). A failed forward extension is not
                                // committed, so in P2 / EMIT b is also bint2, the last good interval (algorithm.hpp:684-699)
};

PGI_HD bool ext_failed(const BiInt &b, uint64_t min_occ) {
    return (uint64_t) b.s < min_occ || b.s <= 0;     // algorithm.hpp:671,690,726 (int64 vs size_t comparison)
}

// `return v` of find_mems_function followed by the `while (x < len)` of find_all_mems and the entry
// checks of the next call (algorithm.hpp:658-666, 745-748).
PGI_HD void lane_return(MemLane &L, uint32_t v, uint32_t min_len, uint64_t n) {
    L.x = v;
    if (L.x >= L.len || L.len - L.x < min_len) { L.x = L.len; L.phase = PH_DONE; return; }
    L.b.k = 0; L.b.kp = 0; L.b.s = (int64_t) n;
    L.j = L.x + min_len - 1; L.phase = PH_P1;
}
PGI_HD void lane_begin(MemLane &L, uint64_t len, uint32_t min_len, uint64_t n) {
    L.len = (uint32_t) (len < PGI_MAX_READ_LEN ? len : PGI_MAX_READ_LEN);
    L.b.k = L.b.kp = 0; L.b.s = 0; L.j = 0;
    lane_return(L, 0, min_len, n);
}

// Advance through the loop heads until the lane needs an extension (LANE_EXTEND; symbol position = L.j, forward iff
// phase == PH_P2), has the MEM {x, j, b.k, b.s} to write (LANE_EMIT; call lane_emitted, then settle again) or is done.
// STRAIGHT-LINE: every transition is a select, taken in the one order in which they can follow each other
//     P3 walked down to x: `return j + 1` -> next find_mems_function call (P1) or done
//     P1 walked down to x: bint2 = bint, j = x + min_len (P2)
//     P2 reached the end of the read: emit
// so the 32 lanes of a warp, which sit in different phases, execute it once together (the switch-in-a-loop form of round 1
// ran once per phase present in the warp: 15 of 32 threads active, a seventh of the kernel's instructions). One call
// passes at most one emit; after an emit the lane is in P3 with j = e >= x + min_len, i.e. it extends next when
// min_len >= 1 (callers that allow min_len = 0 call again, like the host harness does after every emit).
PGI_HD int lane_settle(MemLane &L, uint32_t min_len, uint64_t n) {
    uint32_t ph = L.phase, j = L.j, x = L.x;
    const uint32_t len = L.len;
    // `return v` of find_mems_function, the `while (x < len)` of find_all_mems and the entry checks of the next call
    const bool ret = ph == PH_P3 && j <= x;
    const uint32_t v = j + 1u;
    const bool done = v >= len || len - v < min_len;
    const bool again = ret && !done;
    x = ret ? (done ? len : v) : x;
    ph = ret ? (done ? (uint32_t) PH_DONE : (uint32_t) PH_P1) : ph;
    j = again ? v + min_len - 1u : j;
    L.b.k = again ? 0ull : L.b.k; L.b.kp = again ? 0ull : L.b.kp; L.b.s = again ? (int64_t) n : L.b.s;
    const bool p1_end = ph == PH_P1 && (int32_t) j < (int32_t) x;
    j = p1_end ? x + min_len : j; ph = p1_end ? (uint32_t) PH_P2 : ph;                 // bint2 = bint
    ph = (ph == PH_P2 && j >= len) ? (uint32_t) PH_EMIT : ph;                            // output.push_back({x, e = j, bint2})
    L.phase = ph; L.j = j; L.x = x;
    return ph == PH_EMIT ? LANE_EMIT : ph == PH_DONE ? LANE_DONE : LANE_EXTEND;
}
PGI_HD void lane_emitted(MemLane &L, uint64_t n) {
    L.b.k = 0; L.b.kp = 0; L.b.s = (int64_t) n; L.phase = PH_P3;                       // back = full interval; j = e
}

// Bookkeeping after the extension requested by lane_settle produced t (in the orientation of L.b). Selects only.
// Walking left (P1 / P3) a failure is `return j + 1`: handed to lane_settle as "P3 arrived at x = j"; a success steps to
// j - 1, which may wrap below zero at j = 0 -- lane_settle compares j and x as signed numbers.
PGI_HD void lane_after_ext(MemLane &L, const BiInt &t, uint32_t min_len, uint64_t min_occ, uint64_t n) {
    (void) min_len; (void) n;
    const bool fail = ext_failed(t, min_occ);
    const bool p2 = L.phase == PH_P2;
    const bool commit = !(p2 && fail);                                                  // a failed forward extension is not committed: b stays bint2
    L.b.k = commit ? t.k : L.b.k; L.b.kp = commit ? t.kp : L.b.kp; L.b.s = commit ? t.s : L.b.s;
    const bool left_fail = !p2 && fail;
    L.x = left_fail ? L.j : L.x;
    L.phase = p2 ? (fail ? (uint32_t) PH_EMIT : (uint32_t) PH_P2) : (left_fail ? (uint32_t) PH_P3 : L.phase);
    L.j = p2 ? (fail ? L.j : L.j + 1u) : (fail ? L.j : L.j - 1u);
}

// ------------------------------------------------------------------------------------ seed table
// The first K extensions of every backward walk that starts from the FULL interval {0, 0, n} -- step 1 of
// find_mems_function (algorithm.hpp:666-676) and its step 3 re-scan (:718-735) -- depend only on the K symbols read, so
// their result is tabulated once per index: seed[idx] = the interval after backward-extending {0, 0, n} by the K symbols,
// for every K-mer over {A, C, G, T}, produced by extend_desc itself (so every file-format quirk folded into the
// descriptors is in the table). Interval sizes never grow along a walk when every rank is a true symbol rank (no quirk
// slot: has_quirk == 0, the only case the table is built for), so "the K-th interval passes the min_occ test" implies
// that the K - 1 before it did: one 16-byte lookup then stands for K trips of the kernel, and they are the expensive
// ones (wide intervals: two distinct run blocks per extension). A K-mer with any other byte, or whose entry fails the
// test, walks step by step as before (the failing position matters then).
//
// Window of the read: bytes [lo, lo + K) with lo = (position of the first symbol extended) - K + 1; byte i of the window
// is read position lo + i; extensions run from byte K - 1 down to byte 0. Code of a byte = (byte >> 1) & 3
// (A 0, C 1, T 2, G 3); idx = sum code_i << 2 i.
#define PGI_SEED_MAX_K 15u
struct SeedWord { uint32_t x, y, z, w; };            // k, k', s: low 32 bits in x, y, z; bits 32..39 in the bytes of w
PGI_HD SeedWord seed_pack(const BiInt &b) {
    SeedWord e; e.x = (uint32_t) b.k; e.y = (uint32_t) b.kp; e.z = (uint32_t) (uint64_t) b.s;
    e.w = (uint32_t) ((b.k >> 32) & 0xFFu) | (uint32_t) ((b.kp >> 32) & 0xFFu) << 8 | (uint32_t) (((uint64_t) b.s >> 32) & 0xFFu) << 16;
    return e;
}
PGI_HD BiInt seed_unpack(const SeedWord &e) {
    BiInt b; b.k = (uint64_t) e.x | (uint64_t) (e.w & 0xFFu) << 32; b.kp = (uint64_t) e.y | (uint64_t) ((e.w >> 8) & 0xFFu) << 32;
    b.s = (int64_t) ((uint64_t) e.z | (uint64_t) ((e.w >> 16) & 0xFFu) << 32);
    return b;
}
PGI_HD uint8_t seed_symbol(uint32_t code) { return (uint8_t) ("ACTG"[code & 3u]); }
// The reads of a batch are packed once by a streaming pass (pack_reads_kernel): 2 bits per byte, byte p of the reads
// buffer at bits [2p, 2p + 2) of a little-endian bit stream (one 16-bit unit per aligned 8-byte word), plus one flag per
// read that holds a byte other than 'A', 'C', 'G', 'T' (such a read never uses the table). The index of a window is then
// two 32-bit loads and a funnel shift.
// Codes of the 8 bytes of a word; bad_bytes bit i = byte i is not exactly 'A', 'C', 'G' or 'T'.
PGI_HD uint32_t seed_pack8(uint64_t w, uint32_t &bad_bytes) {
    const uint64_t c = (w >> 1) & 0x0303030303030303ull;
    const uint64_t tf = (c >> 1) & ~c & 0x0101010101010101ull;                         // 'T' (code 2): bit 0 clear, bit 4 set
    const uint64_t e = 0x4040404040404040ull | c << 1 | (tf ^ 0x0101010101010101ull) | tf << 4;   // the byte each code stands for
    uint64_t d = e ^ w;
    d |= d >> 4; d |= d >> 2; d |= d >> 1; d &= 0x0101010101010101ull;
    bad_bytes = (uint32_t) ((d * 0x0102040810204080ull) >> 56);
    uint64_t x = (c | c >> 6) & 0x000F000F000F000Full;
    x = (x | x >> 12) & 0x000000FF000000FFull;
    x = (x | x >> 24) & 0xFFFFull;
    return (uint32_t) x;
}
// Index of the window starting at byte lo of the reads buffer: c0, c1 = the 32-bit units (2 lo) >> 5 and the next one.
PGI_HD uint32_t seed_index_packed(uint32_t c0, uint32_t c1, uint64_t lo, uint32_t K) {
    const uint32_t sh = (uint32_t) (2u * lo) & 31u;
    const uint32_t v = sh ? (c0 >> sh) | (c1 << (32u - sh)) : c0;
    return K >= 16u ? v : v & ((1u << (2u * K)) - 1u);
}
// Does the extension lane_settle just asked for start from the full interval with K symbols in reach?
PGI_HD bool lane_seed_ready(const MemLane &L, uint32_t K, uint64_t n) {
    const bool full = L.b.s == (int64_t) n && (L.b.k | L.b.kp) == 0;
    const bool p1 = L.phase == PH_P1 && L.j + 1u >= L.x + K;                           // symbols j, j - 1, ..., j - K + 1 >= x
    const bool p3 = L.phase == PH_P3 && L.j < L.len && L.j >= L.x + K;                 // ... >= x + 1; pattern[len] is not a read byte
    return full && (p1 || p3);
}
// K successful lane_after_ext steps at once (t = the table entry, known to pass ext_failed). j may wrap below zero like there.
PGI_HD void lane_seed_apply(MemLane &L, const BiInt &t, uint32_t K, uint32_t min_len) {
    (void) min_len;
    L.b = t; L.j -= K;
}
// Entry of K-mer idx: K backward extensions from the full interval; {0, 0, 0} as soon as one comes back empty.
PGI_HD SeedWord seed_entry(const RiDev &ix, const uint32_t *desc_b, uint32_t idx, uint32_t K) {
    BiInt b; b.k = 0; b.kp = 0; b.s = (int64_t) ix.n;
    for (uint32_t i = K; i-- > 0;) {
        b = extend_desc(ix, b, desc_b[seed_symbol(idx >> (2u * i))]);
        if (b.s <= 0) { b.k = 0; b.kp = 0; b.s = 0; break; }
    }
    return seed_pack(b);
}

// ------------------------------------------------------------------------------------ sparse rank
// rank1(i) = number of set positions < i; i >= size counts everything (defined UB).
PGI_HD uint64_t sparse_low(const SparseRankDev &s, uint64_t j) {
    const uint8_t *p = s.low + j * s.low_bytes;
    switch (s.low_bytes) {
    case 1: return *p;
    case 2: return *reinterpret_cast<const uint16_t *>(p);
    case 4: return *reinterpret_cast<const uint32_t *>(p);
    default: return *reinterpret_cast<const uint64_t *>(p);
    }
}
PGI_HD uint64_t sparse_rank1(const SparseRankDev &s, uint64_t i) {
    if (i >= s.size) return s.m;
    const uint64_t b = i >> s.shift;
    uint64_t lo = ld64(s.dir + b), hi = ld64(s.dir + b + 1);
    const uint64_t x = i - (b << s.shift);
    if (hi - lo > 16) {                              // crowded bucket: binary search
        while (lo < hi) { uint64_t mid = (lo + hi) >> 1; if (sparse_low(s, mid) < x) lo = mid + 1; else hi = mid; }
        return lo;
    }
    uint64_t r = lo;
    for (uint64_t j = lo; j < hi; ++j) r += sparse_low(s, j) < x;
    return r;
}

// Predecessor in a sparse set: index j of the last element <= x and its value; false when there is none.
// Replaces sd_vector::predecessor (src/r-index.cpp:1370; blocks_start_pos for the run lookup of locate_encoded).
PGI_HD bool sparse_pred(const SparseRankDev &s, uint64_t x, uint64_t &j, uint64_t &val) {
    if (s.m == 0) return false;
    if (x >= s.size) x = s.size - 1;
    uint64_t b = x >> s.shift;
    const uint64_t r = sparse_rank1(s, x + 1);            // elements <= x
    if (r == 0) return false;
    j = r - 1;
    if (ld64(s.dir + b) > j) {                            // the element sits in an earlier bucket: last b' with dir[b'] <= j
        uint64_t lo = 0, up = b - 1;
        while (lo < up) { const uint64_t mid = (lo + up + 1) >> 1; if (ld64(s.dir + mid) <= j) lo = mid; else up = mid - 1; }
        b = lo;
    }
    val = (b << s.shift) | sparse_low(s, j);
    return true;
}
PGI_HD uint64_t packed_get(const PackedDev &p, uint64_t i) {   // 0 past the end (defined UB)
    if (i >= p.n) return 0;
    const uint64_t bit = i * p.width, w = bit >> 6; const uint32_t sh = (uint32_t) (bit & 63);
    uint64_t v = ld64(p.words + w) >> sh;
    if (sh + p.width > 64) v |= ld64(p.words + w + 1) << (64 - sh);
    return p.width < 64 ? v & ((1ULL << p.width) - 1) : v;
}
// FastLocate::locateNext (phi), src/r-index.cpp:1369-1372.
PGI_HD uint64_t locate_next(const LocDev &l, uint64_t prev) {
    uint64_t j, val;
    if (!sparse_pred(l.last, prev, j, val)) return 0;
    return packed_get(l.samples, packed_get(l.last_to_run, j) + 1) + (prev - val);
}
// Head of FastLocate::locate_encoded (src/r-index.cpp:1308-1329): SA value of BWT row `row` from the sample of its
// logical run plus one phi step per row between the run's first row and `row`.
PGI_HD uint64_t locate_row(const LocDev &l, uint64_t row) {
    uint64_t run = 0, start = 0;
    if (!sparse_pred(l.run_starts, row, run, start)) return 0;
    uint64_t v = packed_get(l.samples, run);
    for (uint64_t p = start; p < row; ++p) v = locate_next(l, v);
    return v;
}

// ------------------------------------------------------------------------------------ tag items
// Plan of one tag query (src/tag_arrays.cpp:857-874): which items to read for the BWT interval [start, end].
struct TagPlan { uint64_t n_runs; uint64_t count; uint64_t first_item; uint64_t skip; };

PGI_HD TagPlan tag_plan(const TagDev &t, uint64_t start, uint64_t end) {
    TagPlan p;
    const uint64_t first = sparse_rank1(t.iv, start + 1);
    const uint64_t last = sparse_rank1(t.iv, end + 1);
    p.n_runs = last - first + 1;                                  // number_of_runs, reported as computed
    p.count = (int64_t) p.n_runs > 0 ? p.n_runs : 0;              // non-positive: the reference loops forever; defined as empty
    const uint64_t dec = first / 10, mv = first % 10;             // encoded_start_every_k_run = 10 (tag_arrays.hpp:120)
    uint64_t sel;
    if (t.starts_regular) sel = dec < t.n_starts ? dec * 10 : t.n_items;
    else sel = dec < t.n_starts ? ld64(t.starts + dec) : t.n_items;   // select beyond the last sample: end of the array (defined UB)
    p.first_item = sel;
    p.skip = mv > 1 ? mv - 1 : 0;                                 // `while (move_tags > 1)`: one run late when first % 10 == 0
    return p;
}

// Item i of the sdsl-compact int_vector<0> (0 past the end), mapped through decode_run_length_compact
// + Position::encode (src/tag_arrays.cpp:47-55,880): identity on node<<11 | rev<<10 | off.
PGI_HD uint64_t tag_item_compact(const TagDev &t, uint64_t i) {
    if (i >= t.n_items) return 0;
    const uint64_t bit = i * t.width, w = bit >> 6; const uint32_t s = (uint32_t) (bit & 63);
    uint64_t v = ld64(t.words + w) >> s;
    if (s + t.width > 64) v |= ld64(t.words + w + 1) << (64 - s);
    if (t.width < 64) v &= (1ULL << t.width) - 1;
    return ((v >> 11) << 11) | (((v >> 10) & 1) << 10) | (v & 0x3FF);
}

// Next ByteCode value of the stream at byte loc (bytes past the end read as 0), advancing loc.
PGI_HD uint64_t tag_bytecode_next(const TagDev &t, uint64_t &loc) {
    uint64_t v = 0; uint32_t sh = 0;
    for (;;) {
        const uint8_t b = loc < t.n_items ? t.stream[loc] : 0; ++loc;
        if (sh < 64) v |= (uint64_t) (b & 0x7F) << sh;
        sh += 7;
        if (!(b & 0x80)) return v;
    }
}
// decode_run (src/tag_arrays.cpp:59-70: off 10 b, rev 1 b, len 9 b, node from bit 20) re-encoded as a position.
PGI_HD uint64_t tag_value_bytecode(uint64_t enc) {
    return ((enc >> 20) << 11) | (((enc >> 10) & 1) << 10) | (enc & 0x3FF);
}

} // namespace pgi
