// pgi_build.cuh — GPU-side construction of the device index from run-length arrays.
//
// What it replaces: in the reference an index reaches the query code through
//   FastLocate(rlbwt) constructor, block part   src/r-index.cpp:778-958   (logical runs -> blocks of 10 runs + cumulatives)
//   serialize_encoded -> load_encoded          src/r-index.cpp:297-376, 406-459
//   compressed tag writers -> load_compressed_tags_compact   src/tag_arrays.cpp:622-654, 766-776, 940-974
// and here through flatten_ri / flatten_tags on the host (pgi_flatten.hpp). For an index that is already resident in
// device memory as a run-length BWT (symbol byte, length) and a run-length tag array (encoded graph position, length)
// -- the output of a BWT construction step -- all of that is a few scans and one pass per bucket on the GPU:
//   merged runs, exclusive scan of their lengths          thrust::reduce_by_key / exclusive_scan
//   first run of every bucket of 2^s BWT positions        vectorised upper_bound over the run starts
//   bucket shift s (fewest crowded buckets rule)          same counts the host flattener uses (pgi_flatten.hpp)
//   one thread per bucket writes its 64-byte block        clipped runs -> codes + end offsets (pgi_layout.h), chunks of
//                                                         crowded buckets behind the bucket array
//   cumulative headers                                    six exclusive scans over the per-bucket symbol counts
//   tag arrays                                            runs split at 511 (tag_arrays.cpp:942-957), items bit-packed at
//                                                         11 + bits(max node) bits, bwt_intervals as the sparse-rank arrays
// The result is the SAME device state pgi_index_create produces from the files the reference's writers would emit for
// those runs (encoded .ri without samples + sdsl-compact tags): tests/test_gpu_parity.py::test_create_from_runs compares
// both indexes query by query and field by field. No host copy of the index exists at any point: config 4
// (n = 7.7e10, 19 GB of run blocks) is built in about a second instead of a minute of host flattening + upload.
#pragma once
#include <thrust/binary_search.h>
#include <thrust/device_ptr.h>
#include <thrust/execution_policy.h>
#include <thrust/functional.h>
#include <thrust/iterator/counting_iterator.h>
#include <thrust/iterator/discard_iterator.h>
#include <thrust/iterator/transform_iterator.h>
#include <thrust/reduce.h>
#include <thrust/scan.h>
#include <thrust/transform.h>
#include <thrust/transform_reduce.h>
#include "pgi_core.cuh"

namespace pgi {
namespace build {

struct CodeOf { __host__ __device__ uint8_t operator()(uint8_t b) const { return b == '\n' ? 0 : b == 'A' ? 1 : b == 'C' ? 2 : b == 'G' ? 3 : b == 'N' ? 4 : b == 'T' ? 5 : 7; } };
struct LenIfCode { uint8_t c; const uint8_t *code; const uint64_t *len; __host__ __device__ uint64_t operator()(uint64_t i) const { return code[i] == c ? len[i] : 0ull; } };
struct LogicalRuns { const uint8_t *code; const uint64_t *len; __host__ __device__ uint64_t operator()(uint64_t i) const { return code[i] == 0 ? len[i] : 1ull; } };
struct BadRun { const uint8_t *code; const uint64_t *len; __host__ __device__ uint64_t operator()(uint64_t i) const { return (code[i] > 5 || len[i] == 0) ? 1ull : 0ull; } };
struct MinusOne { __host__ __device__ uint32_t operator()(uint32_t x) const { return x - 1u; } };
struct Widen { __host__ __device__ uint64_t operator()(uint32_t x) const { return x; } };
struct ZeroLen { const uint64_t *len; __host__ __device__ uint64_t operator()(uint64_t i) const { return len[i] == 0 ? 1ull : 0ull; } };
struct ShiftLeft { unsigned s; __host__ __device__ uint64_t operator()(uint64_t b) const { return b << s; } };

// Clipped runs of bucket b: runs frun[b] .. last, where last = the run holding the bucket's last position.
// (first - 1 for the last bucket when n is a multiple of the bucket size: it starts at position n and holds no run.)
__device__ __forceinline__ uint32_t bucket_last_run(const uint32_t *frun, const uint64_t *starts, uint64_t b, uint64_t nb, unsigned s, uint64_t n_runs, uint64_t n) {
    if (b + 1 >= nb) return (b << s) >= n ? frun[b] - 1 : (uint32_t) (n_runs - 1);
    const uint32_t f = frun[b + 1];
    return starts[f] == ((b + 1) << s) ? f - 1 : f;
}
// End offset (relative to the bucket, clipped to it) of run i.
__device__ __forceinline__ uint32_t clipped_end(const uint64_t *starts, const uint64_t *len, uint32_t i, uint64_t lo, uint64_t hi) {
    const uint64_t e = starts[i] + len[i]; return (uint32_t) ((e < hi ? e : hi) - lo);
}
// Crowded bucket: the largest sub_shift for which every sub-bucket of 2^sub_shift positions holds at most PGI_CHUNK_RUNS
// clipped runs (pgi_flatten.hpp choose_sub_shift).
__device__ unsigned choose_sub_shift_dev(const uint64_t *starts, const uint64_t *len, uint32_t first, uint32_t cnt, uint64_t lo, unsigned S) {
    const uint64_t hi = lo + (1ull << S);
    for (unsigned ss = S - 1;; --ss) {
        bool ok = true; uint32_t i = 0; const uint32_t n_sub = 1u << (S - ss);
        for (uint32_t c = 0; c < n_sub && ok; ++c) {
            const uint32_t hc = (c + 1) << ss; unsigned k = 0;
            while (i < cnt) { ++k; const uint32_t e = clipped_end(starts, len, first + i, lo, hi); if (e > hc) break; ++i; if (e == hc) break; }
            ok = k <= PGI_CHUNK_RUNS;
        }
        if (ok || ss == 0) return ss;
    }
}
// n_chunks == nullptr: only count the crowded buckets (bucket-shift selection). Otherwise n_chunks[b] = chunk blocks of
// bucket b (0 for a normal one) and sub_shift[b] its sub-bucket shift.
__global__ void count_entries_kernel(const uint32_t *frun, const uint64_t *starts, const uint64_t *len, uint64_t nb, unsigned s, uint64_t n_runs, uint64_t n,
                                     uint32_t *n_chunks, uint8_t *sub_shift, unsigned long long *n_crowded) {
    const uint64_t b = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (b >= nb) return;
    const uint32_t first = frun[b], cnt = bucket_last_run(frun, starts, b, nb, s, n_runs, n) - first + 1;
    if (cnt <= PGI_RUNS_PER_BLOCK) { if (n_chunks) { n_chunks[b] = 0; sub_shift[b] = 0; } return; }
    atomicAdd(n_crowded, 1ull);
    if (!n_chunks) return;
    const unsigned ss = choose_sub_shift_dev(starts, len, first, cnt, b << s, s);
    n_chunks[b] = 1u << (s - ss); sub_shift[b] = (uint8_t) ss;
}

// One 64-byte block from up to 12 entries (pgi_flatten.hpp write_block). cum[c] goes into the header as 32-bit partial
// counts here (in-bucket totals for a primary block, counts before the chunk inside its bucket for a chunk block);
// header_kernel turns them into the 40-bit cumulative counts.
__device__ __forceinline__ void put_block(uint32_t *w, const uint32_t cum[6], const uint32_t *code, const uint32_t *end, unsigned ne, uint32_t w0) {
    w[0] = w0; w[1] = cum[1]; w[2] = cum[2]; w[3] = cum[3]; w[4] = cum[5]; w[5] = cum[0]; w[6] = cum[4]; w[7] = 0;
    uint64_t codes = 0;
    for (unsigned i = 0; i < PGI_RUNS_PER_BLOCK; ++i) codes |= (uint64_t) (i < ne ? code[i] : PGI_CODE_UNUSED) << (4 * i);
    w[8] = (uint32_t) (codes >> 32) << 16; w[9] = (uint32_t) codes;
    for (unsigned j = 0; j < 6; ++j) {
        const uint32_t e0 = ne ? end[2 * j < ne ? 2 * j : ne - 1] : 0u, e1 = ne ? end[2 * j + 1 < ne ? 2 * j + 1 : ne - 1] : 0u;
        w[10 + j] = e0 | e1 << 16;
    }
}

__global__ void bucket_kernel(const uint32_t *frun, const uint64_t *starts, const uint64_t *len, const uint8_t *code, uint64_t nb, unsigned s, uint64_t n_runs, uint64_t n,
                              const uint32_t *n_chunks, const uint8_t *sub_shift, const uint64_t *chunk_base, uint32_t *blocks) {
    const uint64_t b = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (b >= nb) return;
    const uint64_t lo = b << s, hi = lo + (1ull << s);
    const uint32_t first = frun[b], last = bucket_last_run(frun, starts, b, nb, s, n_runs, n);
    const uint32_t cnt = last - first + 1;
    uint32_t tot[6] = {0, 0, 0, 0, 0, 0};
    uint32_t *dst = blocks + b * PGI_BLOCK_U32;
    if (cnt <= PGI_RUNS_PER_BLOCK) {
        uint32_t c[PGI_RUNS_PER_BLOCK], e[PGI_RUNS_PER_BLOCK]; uint32_t prev = 0;
        for (uint32_t i = 0; i < cnt; ++i) { c[i] = code[first + i]; e[i] = clipped_end(starts, len, first + i, lo, hi); tot[c[i]] += e[i] - prev; prev = e[i]; }
        put_block(dst, tot, c, e, cnt, 0u);
        return;
    }
    // crowded bucket: directory block + one chunk block per sub-bucket (pad entry + at most 11 runs; pgi_flatten.hpp build_buckets)
    const unsigned ss = sub_shift[b]; const uint32_t n_sub = 1u << (s - ss);
    const uint64_t base = nb + chunk_base[b];
    uint32_t i = 0, at = 0;
    for (uint32_t c = 0; c < n_sub; ++c) {
        const uint32_t lo_c = c << ss, hi_c = (c + 1) << ss;
        uint32_t cc[PGI_RUNS_PER_BLOCK], ce[PGI_RUNS_PER_BLOCK]; unsigned k = 0;
        uint32_t before[6]; for (int q = 0; q < 6; ++q) before[q] = tot[q];
        if (c) { cc[0] = PGI_CODE_UNUSED; ce[0] = lo_c; k = 1; }
        while (i < cnt && at < hi_c && k < PGI_RUNS_PER_BLOCK) {
            const uint32_t ei = clipped_end(starts, len, first + i, lo, hi), e = ei < hi_c ? ei : hi_c;
            cc[k] = code[first + i]; ce[k] = e; ++k; tot[cc[k - 1]] += e - at; at = e;
            if (ei <= hi_c) ++i;
        }
        put_block(blocks + (base + c) * PGI_BLOCK_U32, before, cc, ce, k, 0u);
    }
    put_block(dst, tot, nullptr, nullptr, 0, (uint32_t) base);
    dst[8] &= 0xFFFFu; dst[9] = PGI_CODE_DIRBLOCK; dst[10] = ss;
    for (int j = 11; j < PGI_BLOCK_U32; ++j) dst[j] = 0;
}

// In-bucket total of one header field, read straight out of the primary blocks (scan input).
struct FieldOfBlock { const uint32_t *blocks; unsigned field; __host__ __device__ uint64_t operator()(uint64_t b) const { return blocks[b * PGI_BLOCK_U32 + field]; } };
// field: header word 1..6 (A, C, G, T, NL, N). before[b] = count of that symbol before bucket b.
__global__ void header_kernel(uint32_t *blocks, const uint64_t *before, uint64_t nb, unsigned field, const uint32_t *n_chunks, const uint64_t *chunk_base) {
    const uint64_t b = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (b >= nb) return;
    auto put = [&](uint32_t *w, uint64_t v) {
        w[field] = (uint32_t) v;
        const uint32_t h = (uint32_t) (v >> 32) & 0xFFu;
        if (field == 1) w[7] |= h << 8; else if (field == 2) w[7] |= h << 16; else if (field == 3) w[7] |= h << 24; else if (field == 4) w[8] |= h;
    };
    const uint64_t pre = before[b];
    const uint32_t nc = n_chunks ? n_chunks[b] : 0;
    for (uint32_t c = 0; c < nc; ++c) { uint32_t *w = blocks + (nb + chunk_base[b] + c) * PGI_BLOCK_U32; put(w, pre + w[field]); }
    put(blocks + b * PGI_BLOCK_U32, pre);
}

// Last pass over all blocks: the construction above keeps one count per header word (w1..w4 = A C G T low halves, w5 NL,
// w6 N, high bytes in w7 / w8: scan inputs and outputs are plain words); the layout the queries read is the limb form of
// pgi_layout.h.
__global__ void pack_headers_kernel(uint32_t *blocks, uint64_t n_blocks) {
    const uint64_t b = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (b >= n_blocks) return;
    uint32_t *w = blocks + b * PGI_BLOCK_U32;
    uint64_t cnt[6];
    cnt[0] = w[5]; cnt[4] = w[6];
    cnt[1] = w[1] | (uint64_t) ((w[7] >> 8) & 0xFFu) << 32; cnt[2] = w[2] | (uint64_t) ((w[7] >> 16) & 0xFFu) << 32;
    cnt[3] = w[3] | (uint64_t) (w[7] >> 24) << 32; cnt[5] = w[4] | (uint64_t) (w[8] & 0xFFu) << 32;
    pgi_put_header(w, cnt);
}

// ---- tags
struct Pieces { const uint64_t *len; __host__ __device__ uint64_t operator()(uint64_t i) const { return (len[i] + 510) / 511; } };
struct NodeOf { const uint64_t *val; __host__ __device__ uint64_t operator()(uint64_t i) const { return val[i] >> 11; } };
__global__ void tag_items_kernel(const uint64_t *val, const uint64_t *len, const uint64_t *run_start, const uint64_t *item_off, uint64_t n_runs, uint64_t *item_val, uint64_t *item_pos) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i >= n_runs) return;
    const uint64_t p = (len[i] + 510) / 511, o = item_off[i], s = run_start[i];
    for (uint64_t k = 0; k < p; ++k) { item_val[o + k] = val[i]; item_pos[o + k] = s + 511 * k; }
}
__global__ void pack_items_kernel(const uint64_t *item_val, uint64_t n_items, unsigned width, uint64_t n_words, uint64_t *words) {
    const uint64_t w = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    const uint64_t bit0 = w * 64; uint64_t out = 0;
    const uint64_t mask = width == 64 ? ~0ull : (1ull << width) - 1;
    for (uint64_t i = bit0 / width; i < n_items && i * width < bit0 + 64; ++i) {
        const uint64_t v = item_val[i] & mask, at = i * width;
        out |= at >= bit0 ? v << (at - bit0) : v >> (bit0 - at);
    }
    words[w] = out;
}
__global__ void low_parts_kernel(const uint64_t *pos, uint64_t m, unsigned shift, unsigned low_bytes, uint8_t *low) {
    const uint64_t j = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (j >= m) return;
    const uint64_t v = shift >= 64 ? pos[j] : pos[j] & ((1ull << shift) - 1);
    for (unsigned k = 0; k < low_bytes; ++k) low[j * low_bytes + k] = (uint8_t) (v >> (8 * k));
}
__global__ void iota10_kernel(uint64_t n, uint64_t *out) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i < n) out[i] = 10 * i;
}

} // namespace build
} // namespace pgi
