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
    const uint64_t *blocks;     // 8 words per run block, 64-byte aligned
    const uint64_t *dir;        // block directory
    const uint64_t *q_start;    // quirk tables (n_quirk entries) or nullptr
    const uint64_t *q_val;
    uint64_t n, n_blocks, n_quirk;
    uint32_t dir_shift, has_quirk;
    uint64_t C[8];
};

struct SparseRankDev {
    const uint64_t *dir; const uint8_t *low;
    uint64_t size, m; uint32_t shift, low_bytes;
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
PGI_HD void ld_block(const uint64_t *blk, uint64_t w[8]) {
#if defined(__CUDA_ARCH__)
    const uint4 *q = reinterpret_cast<const uint4 *>(blk);      // four 128-bit loads: one 64-byte block
    uint4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
    w[0] = a.x | ((uint64_t) a.y << 32); w[1] = a.z | ((uint64_t) a.w << 32);
    w[2] = b.x | ((uint64_t) b.y << 32); w[3] = b.z | ((uint64_t) b.w << 32);
    w[4] = c.x | ((uint64_t) c.y << 32); w[5] = c.z | ((uint64_t) c.w << 32);
    w[6] = d.x | ((uint64_t) d.y << 32); w[7] = d.z | ((uint64_t) d.w << 32);
#else
    for (int i = 0; i < 8; ++i) w[i] = blk[i];
#endif
}

// ------------------------------------------------------------------------------------ block locate
// Index of the run block containing BWT position pos (pos <= n): last block whose start <= pos.
// Replaces blocks_start_pos.predecessor(pos) (src/r-index.cpp:594,621): one directory word, and a
// second one only when pos lies behind the first block boundary of its bucket.
PGI_HD uint64_t locate_block(const RiDev &ix, uint64_t pos) {
    const uint64_t b = pos >> ix.dir_shift;
    const uint64_t e = ld64(ix.dir + b);
    uint64_t idx = e & 0xFFFFFFFFull;
    const uint64_t rel = pos - (b << ix.dir_shift);
    if (rel >= (e >> 32)) {
        const uint64_t hi = ld64(ix.dir + b + 1) & 0xFFFFFFFFull;
        if (hi == idx + 1) idx = hi;
        else {                                      // several boundaries in this bucket: search block starts in (idx, hi]
            uint64_t lo = idx + 1, up = hi;          // invariant: start(lo) <= pos
            while (lo < up) {
                uint64_t mid = (lo + up + 1) >> 1;
                uint64_t st = ld64(ix.blocks + mid * PGI_BLOCK_WORDS) & 0xFFFFFFFFFFull;
                if (st <= pos) lo = mid; else up = mid - 1;
            }
            idx = lo;
        }
    }
    return idx;
}

// Symbol counts before a block, per nuc code, from its header words.
PGI_HD void block_cums(const uint64_t w[8], uint64_t cum[6]) {
    cum[1] = (w[0] >> 40) | ((w[1] & 0xFFFFull) << 24);
    cum[2] = (w[1] >> 16) & 0xFFFFFFFFFFull;
    cum[3] = w[2] & 0xFFFFFFFFFFull;
    cum[5] = (w[2] >> 40) | ((w[3] & 0xFFFFull) << 24);
    cum[0] = (w[3] >> 16) & 0xFFFFFFFFull;
    cum[4] = (w[3] >> 48) | ((w[4] & 0xFFFFull) << 16);
}

// 16-bit run entry i (0..14) of a block.
PGI_HD uint32_t block_entry(const uint64_t w[8], int i) {
    const int bit = 16 + 16 * i;                    // relative to word 4
    return (uint32_t) (w[4 + (bit >> 6)] >> (bit & 63)) & 0xFFFFu;
}

// In-block counts of the six codes among the first rel1 and the first rel2 positions of a block.
// Packed accumulators: codes 0..2 in lo, 3..5 in hi, 21 bits per field (a block covers < 2^17 positions).
struct Packed6 { uint64_t lo, hi; };
PGI_HD void block_scan2(const uint64_t w[8], uint64_t rel1, uint64_t rel2, Packed6 &c1, Packed6 &c2) {
    c1.lo = c1.hi = c2.lo = c2.hi = 0;
    // a block covers at most 15 * 8191 positions; clamp so the 32-bit arithmetic below is exact
    uint32_t r1 = rel1 > 0x3FFFFull ? 0x3FFFFu : (uint32_t) rel1;
    uint32_t r2 = rel2 > 0x3FFFFull ? 0x3FFFFu : (uint32_t) rel2;
#pragma unroll
    for (int i = 0; i < PGI_RUNS_PER_BLOCK; ++i) {
        const uint32_t e = block_entry(w, i);
        const uint32_t len = e >> 3, code = e & 7;
        const uint32_t t1 = len < r1 ? len : r1, t2 = len < r2 ? len : r2;
        r1 -= t1; r2 -= t2;
        const bool up = code >= 3;
        const uint32_t sh = (code - (up ? 3u : 0u)) * 21u;
        const uint64_t v1 = (uint64_t) t1 << sh, v2 = (uint64_t) t2 << sh;
        if (up) { c1.hi += v1; c2.hi += v2; } else { c1.lo += v1; c2.lo += v2; }
    }
}
PGI_HD uint64_t packed_get(const Packed6 &p, int c) {
    return ((c >= 3 ? p.hi : p.lo) >> (21 * (c >= 3 ? c - 3 : c))) & 0x1FFFFFull;
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

// True ranks of the six nuc codes at pos1 <= n and pos2 <= n (either order).
// Both positions usually share a block once the interval is short: one 64-byte load serves both.
PGI_HD void rank_pair(const RiDev &ix, uint64_t pos1, uint64_t pos2, uint64_t tr1[6], uint64_t tr2[6]) {
    const uint64_t b1 = locate_block(ix, pos1), b2 = locate_block(ix, pos2);
    uint64_t w[8], cum[6]; Packed6 c1, c2;
    ld_block(ix.blocks + b1 * PGI_BLOCK_WORDS, w);
    const uint64_t st1 = w[0] & 0xFFFFFFFFFFull;
    if (b1 == b2) {
        block_scan2(w, pos1 - st1, pos2 - st1, c1, c2);
        block_cums(w, cum);
#pragma unroll
        for (int c = 0; c < 6; ++c) { tr1[c] = cum[c] + packed_get(c1, c); tr2[c] = cum[c] + packed_get(c2, c); }
    } else {
        uint64_t v[8];
        ld_block(ix.blocks + b2 * PGI_BLOCK_WORDS, v);
        Packed6 d;
        block_scan2(w, pos1 - st1, 0, c1, d);
        block_cums(w, cum);
#pragma unroll
        for (int c = 0; c < 6; ++c) tr1[c] = cum[c] + packed_get(c1, c);
        const uint64_t st2 = v[0] & 0xFFFFFFFFFFull;
        block_scan2(v, pos2 - st2, 0, c2, d);
        block_cums(v, cum);
#pragma unroll
        for (int c = 0; c < 6; ++c) tr2[c] = cum[c] + packed_get(c2, c);
    }
}

// ------------------------------------------------------------------------------------ FMD extension
struct BiInt { uint64_t k, kp; int64_t s; };      // FastLocate::bi_interval {forward, reverse, size}

// One backward extension with the per-byte descriptor d (desc_b[a]); forward extension = the same on
// the swapped interval with desc_f[a] (src/r-index.cpp:758-764). All arithmetic mod 2^64, like the
// reference's size_t. Positions beyond n read as n (defined UB).
PGI_HD BiInt extend_desc(const RiDev &ix, BiInt in, uint32_t d) {
    uint64_t p2 = in.k + (uint64_t) in.s, p1 = in.k;
    if (p2 > ix.n) p2 = ix.n;
    if (p1 > ix.n) p1 = ix.n;
    const uint64_t b1 = locate_block(ix, p1), b2 = locate_block(ix, p2);
    uint64_t w[8], v[8];
    ld_block(ix.blocks + b1 * PGI_BLOCK_WORDS, w);
    const bool same = b1 == b2;
    if (!same) ld_block(ix.blocks + b2 * PGI_BLOCK_WORDS, v);     // both block loads are in flight together
    const uint32_t ms = (d >> 21) & 7;
    uint64_t kp = in.kp, x = 0, y = 0;
    {
        const uint64_t st = w[0] & 0xFFFFFFFFFFull;
        Packed6 c1, c2; uint64_t cum[6];
        block_scan2(w, p1 - st, same ? p2 - st : 0, c1, c2);
        block_cums(w, cum);
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            const uint64_t wt = (d >> (3 * c)) & 7;
            const uint64_t r1 = cum[c] + packed_get(c1, c);
            kp -= wt * r1;
            if (ms == (uint32_t) c) y = r1;
            if (same) {
                const uint64_t r2 = cum[c] + packed_get(c2, c);
                kp += wt * r2;
                if (ms == (uint32_t) c) x = r2;
            }
        }
    }
    if (!same) {
        const uint64_t st = v[0] & 0xFFFFFFFFFFull;
        Packed6 c2, dd; uint64_t cum[6];
        block_scan2(v, p2 - st, 0, c2, dd);
        block_cums(v, cum);
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            const uint64_t r2 = cum[c] + packed_get(c2, c);
            kp += (uint64_t) ((d >> (3 * c)) & 7) * r2;
            if (ms == (uint32_t) c) x = r2;
        }
    }
    if (ix.has_quirk) {
        const uint32_t wq = (d >> 18) & 7;
        if (wq || ms == PGI_SRC_QUIRK) {
            const uint64_t q2 = quirk_at(ix, p2), q1 = quirk_at(ix, p1);
            kp += (uint64_t) wq * (q2 - q1);
            if (ms == PGI_SRC_QUIRK) { x = q2; y = q1; }
        }
    }
    BiInt out;
    if (y >= x) { out.k = 0; out.kp = 0; out.s = 0; return out; }
    out.k = y + ix.C[(d >> 24) & 7]; out.kp = kp; out.s = (int64_t) (x - y);
    return out;
}

// ------------------------------------------------------------------------------------ MEM state machine
// One lane walks one read through find_all_mems as a loop whose every trip performs exactly one FMD
// extension (so the 32 lanes of a warp stay convergent although they are in different phases).
enum { PH_P1 = 1, PH_P2 = 2, PH_P3 = 3, PH_DONE = 0 };

struct MemLane {
    uint64_t len, x, j, e;
    BiInt b;                    // interval being extended (bint in P1/P2, back in P3)
    uint64_t k2; int64_t s2;    // bint2.forward / bint2.size
    uint32_t phase, emit;       // emit = 1: a MEM {x, e, k2, s2} is ready to be written
};

PGI_HD bool ext_failed(const BiInt &b, uint64_t min_occ) {
    return (uint64_t) b.s < min_occ || b.s <= 0;     // algorithm.hpp:671,690,726 (int64 vs size_t comparison)
}

// `return v` of find_mems_function followed by the `while (x < len)` of find_all_mems and the entry
// checks of the next call (algorithm.hpp:658-666, 745-748).
PGI_HD void lane_return(MemLane &L, uint64_t v, uint64_t min_len, uint64_t n) {
    L.x = v;
    if (L.x >= L.len) { L.phase = PH_DONE; return; }
    if (L.len - L.x < min_len) { L.x = L.len; L.phase = PH_DONE; return; }
    L.b.k = 0; L.b.kp = 0; L.b.s = (int64_t) n;
    L.j = L.x + min_len - 1; L.phase = PH_P1;
}
PGI_HD void lane_begin(MemLane &L, uint64_t len, uint64_t min_len, uint64_t n) {
    L.len = len; L.emit = 0; L.e = 0; L.k2 = 0; L.s2 = 0;
    L.b.k = L.b.kp = 0; L.b.s = 0; L.j = 0;
    lane_return(L, 0, min_len, n);
}

// Advance through loop heads until the lane needs an extension (returns true; symbol position = L.j,
// forward iff phase == PH_P2), has a MEM to emit (L.emit set; call again afterwards) or is done.
PGI_HD bool lane_settle(MemLane &L, uint64_t min_len, uint64_t n) {
    for (;;) {
        if (L.emit) return false;
        switch (L.phase) {
        case PH_P1:
            if ((int64_t) L.j >= (int64_t) L.x) return true;
            L.k2 = L.b.k; L.s2 = L.b.s; L.j = L.x + min_len; L.phase = PH_P2;        // bint2 = bint
            break;
        case PH_P2:
            if (L.j < L.len) return true;
            L.e = L.j; L.emit = 1;                                                     // output.push_back
            L.b.k = 0; L.b.kp = 0; L.b.s = (int64_t) n; L.phase = PH_P3;               // back = full; j = e
            return false;
        case PH_P3:
            if (L.j > L.x) return true;
            lane_return(L, L.j + 1, min_len, n);
            break;
        default:
            return false;
        }
    }
}

// Bookkeeping after the extension requested by lane_settle produced L.b.
PGI_HD void lane_after_ext(MemLane &L, uint64_t min_len, uint64_t min_occ, uint64_t n) {
    const bool fail = ext_failed(L.b, min_occ);
    switch (L.phase) {
    case PH_P1:
        if (fail) lane_return(L, L.j + 1, min_len, n);
        else if (L.j == 0) { L.k2 = L.b.k; L.s2 = L.b.s; L.j = L.x + min_len; L.phase = PH_P2; }
        else --L.j;
        break;
    case PH_P2:
        if (fail) { L.e = L.j; L.emit = 1; L.b.k = 0; L.b.kp = 0; L.b.s = (int64_t) n; L.phase = PH_P3; }
        else { L.k2 = L.b.k; L.s2 = L.b.s; ++L.j; }
        break;
    case PH_P3:
        if (fail) lane_return(L, L.j + 1, min_len, n);
        else --L.j;
        break;
    default: break;
    }
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
