// pgi_tag_kernels.cuh — tag-array lookup kernels (K2a/K2b/K3/K4 of pgi_kernels.cuh's table) and the
// random-access microbenchmark. Semantics: TagArray::query_compressed_compact / query_compressed,
// src/tag_arrays.cpp:780-890 (rank on bwt_intervals, sampled select, decode, sort, unique).
#pragma once
#include "pgi_kernels.cuh"

namespace pgi {

// The number of MEMs (nm) of a step is known on the device only (state word written by mem_finalize_kernel): every kernel
// below is launched over the CAPACITY of the MEM buffers and takes nm from *n_dev, so that one find_mems step is a single
// stream-ordered sequence without host round trips (pgi_capi.cu batch_run).
//
// K2a. One thread per MEM: the two ranks and the select; sizes only.
// Intervals come either from ordered MEM records (mems != nullptr: [bwt_start, bwt_start + size - 1])
// or from explicit start/end arrays. cnt[] is zeroed beforehand (entries >= nm stay 0 for the scan).
__global__ void tag_plan_kernel(TagDev t, uint64_t cap, const uint64_t *n_dev, const uint4 *mems, const uint64_t *start, const uint64_t *end,
                                uint64_t *n_runs, uint64_t *cnt, uint64_t *first_item, uint32_t *skip) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i >= cap || i >= *n_dev) return;
    uint64_t s, e;
    if (mems) {
        const uint4 r = __ldg(mems + 2 * i + 1);
        s = r.x | ((uint64_t) r.y << 32);
        e = s + (r.z | ((uint64_t) r.w << 32)) - 1;
    } else { s = start[i]; e = end[i]; }
    if (mems && e + 1 == s) { n_runs[i] = 0; cnt[i] = 0; first_item[i] = 0; skip[i] = 0; return; }   // size 0: a query_tags read without a match is not queried (src/query_tags.cpp:97-100); MEMs always have size > 0
    const TagPlan p = tag_plan(t, s, e);
    n_runs[i] = p.n_runs; cnt[i] = p.count; first_item[i] = p.first_item; skip[i] = (uint32_t) p.skip;
}

// Tag-stage state words (device): [0] number of long lists queued (keeps counting past large_cap), [1] set when a list
// did not fit the raw buffer (raw_cap).
// K2b (short lists). One thread per MEM: decode <= PGI_TAG_SMALL runs, insertion sort, unique; the unique
// values go to raw[raw_off[m] ..) and their number to ucnt[m]. Longer lists are queued for the warp kernels.
__global__ void tag_small_kernel(TagDev t, uint64_t cap, const uint64_t *n_dev, const uint64_t *cnt, const uint64_t *raw_off, const uint64_t *first_item,
                                 const uint32_t *skip, uint64_t *raw, uint64_t raw_cap, uint64_t *ucnt, uint64_t *large_list, uint64_t large_cap,
                                 unsigned long long *tag_state) {
    const uint64_t m = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (m >= cap || m >= *n_dev) return;
    const uint64_t c = cnt[m];
    if (c == 0) { ucnt[m] = 0; return; }
    if (raw_off[m] + c > raw_cap) { tag_state[1] = 1ull; ucnt[m] = 0; return; }
    if (c > PGI_TAG_SMALL) { const unsigned long long q = atomicAdd(tag_state, 1ull); if (q < large_cap) large_list[q] = m; ucnt[m] = 0; return; }
    uint64_t v[PGI_TAG_SMALL];
    uint32_t k = 0;                                     // sorted unique prefix v[0..k)
    uint64_t loc = first_item[m]; const uint32_t sk = skip[m];
    if (t.layout != 0) for (uint32_t i = 0; i < sk; ++i) (void) tag_bytecode_next(t, loc);
    else loc += sk;
    for (uint32_t i = 0; i < (uint32_t) c; ++i) {
        const uint64_t x = t.layout == 0 ? tag_item_compact(t, loc++) : tag_value_bytecode(tag_bytecode_next(t, loc));
        uint32_t p = k;                                  // insert keeping v sorted, dropping duplicates
        while (p > 0 && v[p - 1] > x) --p;
        if (p > 0 && v[p - 1] == x) continue;
        for (uint32_t q = k; q > p; --q) v[q] = v[q - 1];
        v[p] = x; ++k;
    }
    uint64_t *dst = raw + raw_off[m];
    for (uint32_t i = 0; i < k; ++i) dst[i] = v[i];
    ucnt[m] = k;
}

// K2b (long lists). One warp per queued MEM (grid-stride over the queue, whose length is on the device): decode its runs
// into raw (lane-strided for the bit-packed layout; the ByteCode stream is inherently serial and is walked by lane 0).
// Segments of queue slots that stay unused are empty (seg arrays are zeroed beforehand).
__device__ __forceinline__ uint64_t queued(const unsigned long long *tag_state, uint64_t large_cap) { const uint64_t q = tag_state[0]; return q < large_cap ? q : large_cap; }
__global__ void tag_large_decode_kernel(TagDev t, const unsigned long long *tag_state, uint64_t large_cap, const uint64_t *large_list, const uint64_t *cnt,
                                        const uint64_t *raw_off, const uint64_t *first_item, const uint32_t *skip,
                                        uint64_t *raw, uint64_t *seg_begin, uint64_t *seg_end) {
    const unsigned lane = threadIdx.x & 31u;
    const uint64_t n_large = queued(tag_state, large_cap), n_warps = ((uint64_t) gridDim.x * blockDim.x) >> 5;
    for (uint64_t w = (blockIdx.x * (uint64_t) blockDim.x + threadIdx.x) >> 5; w < n_large; w += n_warps) {
        const uint64_t m = large_list[w], c = cnt[m], o = raw_off[m];
        if (lane == 0) { seg_begin[w] = o; seg_end[w] = o + c; }
        if (t.layout == 0) {
            const uint64_t f = first_item[m] + skip[m];
            for (uint64_t i = lane; i < c; i += 32) raw[o + i] = tag_item_compact(t, f + i);
        } else if (lane == 0) {
            uint64_t loc = first_item[m];
            for (uint32_t i = 0, sk = skip[m]; i < sk; ++i) (void) tag_bytecode_next(t, loc);
            for (uint64_t i = 0; i < c; ++i) raw[o + i] = tag_value_bytecode(tag_bytecode_next(t, loc));
        }
    }
}

// K3 (long lists). One warp per queued MEM: unique of the sorted segment in `sorted`, written back to raw.
__global__ void tag_large_unique_kernel(const unsigned long long *tag_state, uint64_t large_cap, const uint64_t *large_list, const uint64_t *cnt, const uint64_t *raw_off,
                                        const uint64_t *sorted, uint64_t *raw, uint64_t *ucnt) {
    const unsigned lane = threadIdx.x & 31u;
    const uint64_t n_large = queued(tag_state, large_cap), n_warps = ((uint64_t) gridDim.x * blockDim.x) >> 5;
    for (uint64_t w = (blockIdx.x * (uint64_t) blockDim.x + threadIdx.x) >> 5; w < n_large; w += n_warps) {
        const uint64_t m = large_list[w], c = cnt[m], o = raw_off[m];
        uint64_t k = 0;
        for (uint64_t b = 0; b < c; b += 32) {
            const uint64_t i = b + lane;
            bool keep = false; uint64_t x = 0;
            if (i < c) { x = sorted[o + i]; keep = i == 0 || sorted[o + i - 1] != x; }
            const unsigned mask = __ballot_sync(0xFFFFFFFFu, keep);
            if (keep) raw[o + k + __popc(mask & ((1u << lane) - 1u))] = x;
            k += __popc(mask);
        }
        if (lane == 0) ucnt[m] = k;
    }
}

// K4. Compaction: unique lists -> tags[tag_off[m] ..). Short lists: one thread per MEM.
__global__ void tag_copy_kernel(uint64_t cap, const uint64_t *n_dev, const uint64_t *cnt, const uint64_t *raw_off, const uint64_t *tag_off,
                                const uint64_t *raw, uint64_t *tags) {
    const uint64_t m = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (m >= cap || m >= *n_dev || cnt[m] > PGI_TAG_SMALL) return;
    const uint64_t a = tag_off[m], k = tag_off[m + 1] - a; const uint64_t *src = raw + raw_off[m];
    for (uint64_t i = 0; i < k; ++i) tags[a + i] = src[i];
}
__global__ void tag_large_copy_kernel(const unsigned long long *tag_state, uint64_t large_cap, const uint64_t *large_list, const uint64_t *raw_off, const uint64_t *tag_off,
                                      const uint64_t *raw, uint64_t *tags) {
    const unsigned lane = threadIdx.x & 31u;
    const uint64_t n_large = queued(tag_state, large_cap), n_warps = ((uint64_t) gridDim.x * blockDim.x) >> 5;
    for (uint64_t w = (blockIdx.x * (uint64_t) blockDim.x + threadIdx.x) >> 5; w < n_large; w += n_warps) {
        const uint64_t m = large_list[w], a = tag_off[m], k = tag_off[m + 1] - a, o = raw_off[m];
        for (uint64_t i = lane; i < k; i += 32) tags[a + i] = raw[o + i];
    }
}

// M1. Random-access roofline: every thread walks `chain` dependent hops through a buffer of `n_slots`
// slots of BYTES bytes, reading the whole slot each hop (BYTES = 32: one sector; 64: one run block).
template <int BYTES>
__global__ void gather_kernel(const uint4 *buf, uint64_t n_slots, uint64_t n_threads, int chain, unsigned long long *sink) {
    const uint64_t tid = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (tid >= n_threads) return;
    uint64_t x = tid * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull, acc = 0;
    for (int h = 0; h < chain; ++h) {
        x ^= x >> 29; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 32;
        const uint64_t slot = x % n_slots;
        const uint4 *p = buf + slot * (BYTES / 16);
        uint64_t s = 0;
#pragma unroll
        for (int q = 0; q < BYTES / 16; ++q) { const uint4 v = __ldg(p + q); s += v.x + v.y + v.z + v.w; }
        acc += s; x += s;                                // next address depends on the loaded data
    }
    if (acc == 0x1234567ull) atomicAdd(sink, 1ull);
}

// M1, independent form (chain == 1): every thread issues `per_thread` random loads whose addresses do not depend on
// loaded data (4 in flight per thread), over n_slots slots of BYTES bytes. This is the throughput ceiling of random
// sector access for a given footprint; the chained form above adds the dependent-latency component.
template <int BYTES>
__global__ void gather_indep_kernel(const uint4 *buf, uint32_t n_slots, uint32_t per_thread, unsigned long long *sink) {
    const uint64_t tid = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    uint64_t x = tid * (uint64_t) per_thread * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull; uint32_t acc = 0;   // a distinct counter per access
    for (uint32_t i = 0; i < per_thread; i += 4) {
        uint32_t slot[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            x += 0x9E3779B97F4A7C15ull; uint64_t z = x; z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27; z *= 0x94D049BB133111EBull; z ^= z >> 31;
            slot[u] = (uint32_t) (((z >> 32) * (uint64_t) n_slots) >> 32);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint4 *p = buf + (uint64_t) slot[u] * (BYTES / 16);
#pragma unroll
            for (int q = 0; q < BYTES / 16; ++q) { const uint4 v = __ldg(p + q); acc += v.x ^ v.y ^ v.z ^ v.w; }
        }
    }
    if (acc == 0x12345u) atomicAdd(sink, 1ull);
}

// M1, cooperative form (chain == 0): the access pattern of fmd_mem_kernel's block fetch without any of its arithmetic. A TEAM
// of BYTES / 16 adjacent lanes loads one random slot (16 bytes per lane: one request of the warp covers 32 / team slots),
// eight independent slots in flight per lane, at full occupancy: about 260 KB of requests in flight per SM, an order of
// magnitude more than latency x bandwidth needs. Nothing depends on loaded data. This is the CEILING for random BYTES-byte
// bursts over a footprint of n_slots x BYTES (the figure a gather-bound kernel is measured against).
template <int BYTES>
__global__ void __launch_bounds__(256) gather_team_kernel(const uint4 *buf, uint32_t n_slots, uint32_t rounds, unsigned long long *sink) {
    constexpr unsigned TEAM = BYTES / 16;
    const uint64_t tid = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    const uint64_t team = tid / TEAM; const unsigned part = (unsigned) (tid % TEAM);
    uint64_t x = team * (uint64_t) rounds * 8u * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull; uint32_t acc = 0;
    for (uint32_t r = 0; r < rounds; ++r) {
        uint4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x += 0x9E3779B97F4A7C15ull; uint64_t z = x; z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27; z *= 0x94D049BB133111EBull; z ^= z >> 31;
            const uint32_t slot = (uint32_t) (((z >> 32) * (uint64_t) n_slots) >> 32);
            v[u] = __ldg(buf + (uint64_t) slot * TEAM + part);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345u) atomicAdd(sink, 1ull);
}

// ------------------------------------------------------------------------------------ locate (toehold + phi)
// FastLocate::locate_encoded (src/r-index.cpp:1300-1343), one thread per BWT range [first, second]: SA value of the first
// row from the sample of its run, then one phi step (locateNext, :1369-1372) per further row -- an inherently serial
// chain of four dependent random accesses per row; parallelism comes from the number of ranges in flight.
// sa[row_off[i] + t] = SA value of row first + t; ids = the same divided by max_length (seqId, r-index.hpp:429).
__global__ void locate_kernel(LocDev l, uint64_t n, const uint64_t *first, const uint64_t *second, const uint64_t *row_off,
                              uint64_t *sa, uint64_t *ids) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t f = first[i], e = second[i];
    if (e < f) return;
    uint64_t o = row_off[i];
    uint64_t v = locate_row(l, f);
    sa[o] = v; ids[o] = v / l.max_length;
    for (uint64_t r = f; r < e; ++r) { v = locate_next(l, v); ++o; sa[o] = v; ids[o] = v / l.max_length; }
}
// std::unique over the sorted ids of every range, in place; ucnt[i] = number kept.
__global__ void locate_unique_kernel(uint64_t n, const uint64_t *row_off, uint64_t *ids, uint64_t *ucnt) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t a = row_off[i], b = row_off[i + 1];
    uint64_t k = a;
    for (uint64_t p = a; p < b; ++p) { const uint64_t x = ids[p]; if (k == a || ids[k - 1] != x) ids[k++] = x; }
    ucnt[i] = k - a;
}
__global__ void locate_copy_kernel(uint64_t n, const uint64_t *row_off, const uint64_t *id_off, const uint64_t *ids, uint64_t *out) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t a = row_off[i], d = id_off[i], c = id_off[i + 1] - d;
    for (uint64_t t = 0; t < c; ++t) out[d + t] = ids[a + t];
}

} // namespace pgi
