// pgi_kernels.cuh — the CUDA kernels of the find_mems hot path (sm_100a). Integer / bit work over
// random-access HBM traffic: no tensor cores, nothing here is a dense contraction.
//
//   fmd_mem_kernel      K1  find_all_mems for a batch of reads      (algorithm.hpp:653-757)
//   mem_scatter_kernel  K4  MEMs into read order
//   tag_plan_kernel     K2a rank/select part of query_compressed*   (tag_arrays.cpp:857-874)
//   tag_small_kernel    K2b decode + sort + unique, short lists
//   tag_large_*         K2b/K3 the same for long lists, one warp per MEM
//   tag_copy_kernel     K4  compaction of the unique lists
//   fmd_extend_kernel / rank_kernel : batched single operations (parity tests, facade calls)
//   gather_kernel       M1  random-access roofline microbenchmark
#pragma once
#include <cuda_runtime.h>
#include "pgi_core.cuh"

namespace pgi {

#define PGI_TAG_SMALL 32u          // lists up to this many runs are handled by one thread

struct MemArgs {
    RiDev ix;
    const uint32_t *desc;          // desc_b[256], desc_f[256], then the 512 x 4-word scan tables of those descriptors (PGI_DESC_WORDS)
    const uint64_t *reads64;       // read bytes viewed as aligned u64 words (buffer padded to 8 bytes)
    const uint64_t *read_off;      // n_reads + 1 byte offsets
    uint64_t n_reads, min_len, min_occ;
    unsigned long long *cursor;    // next unassigned read
    unsigned long long *mem_count; // MEMs appended so far (keeps counting past mem_cap)
    unsigned long long *ext_count; // FMD extensions of the reference algorithm (a seed lookup counts as its seed_k extensions)
    unsigned long long *seed_count;// seed lookups that replaced seed_k extensions each
    unsigned long long *bad_off;   // set when a read's offsets are not ascending / leave the reads buffer (the step then fails with PGI_EINVAL)
    uint64_t reads_bytes;          // bytes of the reads buffer
    unsigned long long *blk_count; // 64-byte run blocks fetched (bucket blocks + chunk blocks): what the kernel moves from the index
    uint64_t mem_cap;
    uint4 *mem_buf;                // 2 x uint4 per appended MEM {start, end, bwt_start, size}
    uint64_t *mem_key;             // read id << 32 | index of the MEM within its read
    uint64_t *read_nmems;          // n_reads (+1) MEM counts per read
    const uint32_t *codes32;       // 2-bit codes of the reads buffer (pack_reads_kernel), viewed as 32-bit units
    const uint32_t *read_flags;    // bit r: read r holds a byte other than A / C / G / T (no seed lookups for it)
    const uint4 *seed;             // seed table (pgi_core.cuh): 4^seed_k entries of 16 bytes, or nullptr
    uint32_t seed_k, pad;          // 0: no table (OP 1, indexes with a quirk slot, PGI_SEEDK=0)
    // MEM slots. One global counter costs a same-address atomicAdd per MEM (3.8 M per 2 M reads: one every 2.3 ns), and how
    // fast the L2 retires those differs between GPUs and moments: 8.8 or 10.9 ms for the same launch.
    // With part_cap != 0 the append buffer is cut into PGI_SLOT_PARTS partitions of part_cap slots, each with its own
    // counter on its own 128-byte line; a warp appends to partition (global warp id mod PGI_SLOT_PARTS). The scatter pass
    // walks the filled prefix of every partition, so the holes behind them never matter.
    unsigned long long *slot_ctr;  // PGI_SLOT_PARTS counters, 16 u64 apart (keep counting past part_cap)
    uint64_t part_cap;             // 0: the single counter mem_count
};
#define PGI_SLOT_PARTS 256u

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// K1. One lane = one read at a time; every trip of the loop performs exactly one FMD extension per
// active lane, so a warp stays convergent while its 32 reads sit in different phases of
// find_mems_function. Lanes that finish a read pull the next one from a warp-local queue that is
// refilled 32 reads at a time from a global cursor (persistent grid: no tail of idle lanes inside a warp).
// Read bytes and MEM records are streamed (evict-first) so that they do not push the index out of L2.
//
// Block fetch (MODE): every lane needs its own one or two 64-byte run blocks per extension.
//   MODE 0  each lane loads its own blocks (4 x LDG.128 per block). 32 lanes = 32 distinct cache lines per load
//           instruction: the L1 data pipe pays one wavefront per line, 4 x 32 per block and warp (ncu r01e: L1 wavefront
//           throughput 79 % of peak, the limiter).
//   MODE 1  cooperative: a quad of lanes fetches one block (4 x 16 B contiguous), so one instruction of the warp covers
//           8 blocks = at most 8 lines; 4 rounds serve the 32 lanes. The 16-byte parts are copied global -> shared with
//           cp.async (no registers, L1 bypassed) into a per-warp staging area, XOR-swizzled so that both the quad-wise
//           writes and the per-lane 4 x LDS.128 reads are free of bank conflicts. 4x fewer L1 wavefronts per block.
//           (LDG.128 + STS.128 instead of cp.async was measured as well: 17.8 ms against 16.6 ms per 2 M reads.)
#define PGI_NOBLK 0xFFFFFFFFu
// Bounds-checking build (make -C csrc bounds -> libpgi_bounds.so; PGI_LIB selects it in the tests): every block / chunk
// index, staging slot and MEM slot is checked before use and a violation traps (compute-sanitizer is closed on this pool).
#ifdef PGI_BOUNDS
#define PGI_CHECK(cond) do { if (!(cond)) { printf("PGI_BOUNDS violated: %s (line %d)\n", #cond, __LINE__); __trap(); } } while (0)
#else
#define PGI_CHECK(cond) do { } while (0)
#endif
// 16-byte global -> shared copy, skipped (predicated off inside the asm: no branch, no address re-derivation) when blk == PGI_NOBLK.
__device__ __forceinline__ void cp_async16_if(uint32_t saddr, const void *g, uint32_t blk) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0xffffffff;\n\t@p cp.async.cg.shared.global [%0], [%1], 16;\n\t}" ::"r"(saddr), "l"(g), "r"(blk) : "memory");
}
// The same with the address formed inside: 16 bytes at base + 64 * blk (one IMAD.WIDE; MODE 2).
__device__ __forceinline__ void cp_async16_blk(uint32_t saddr, const char *base, uint32_t blk) {
    asm volatile("{\n\t.reg .pred p;\n\t.reg .u64 a;\n\tsetp.ne.u32 p, %2, 0xffffffff;\n\tmad.wide.u32 a, %2, 64, %1;\n\t@p cp.async.cg.shared.global [%0], [a], 16;\n\t}" ::"r"(saddr), "l"(base), "r"(blk) : "memory");
}
// MODE 4: the same copy with an L2 evict-first policy: run blocks and seed entries are random accesses without reuse, and at
// 5 TB/s they turn the whole L2 over every 25 us, taking the batch's 2-bit codes (75 MB, written by pack_reads_kernel just
// before, read again window by window) with them. Measured: 7.98 ms against 8.04 ms (profiles/r03e; asking the code loads
// for evict-last on top changed nothing).
__device__ __forceinline__ uint64_t l2_evict_first_policy() { uint64_t p; asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ void cp_async16_blk_ef(uint32_t saddr, const char *base, uint32_t blk, uint64_t pol) {
    asm volatile("{\n\t.reg .pred p;\n\t.reg .u64 a;\n\tsetp.ne.u32 p, %2, 0xffffffff;\n\tmad.wide.u32 a, %2, 64, %1;\n\t@p cp.async.cg.shared.global.L2::cache_hint [%0], [a], 16, %3;\n\t}" ::"r"(saddr), "l"(base), "r"(blk), "l"(pol) : "memory");
}
__device__ __forceinline__ void cp_async16_if_ef(uint32_t saddr, const void *g, uint32_t blk, uint64_t pol) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0xffffffff;\n\t@p cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %3;\n\t}" ::"r"(saddr), "l"(g), "r"(blk), "l"(pol) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Words of the lane's own staged block (slot = lane) of position `which`.
// MODE 2: instead of XOR-swizzling the four parts inside a slot, slot L starts 16 (L >> 1) bytes late (16 bytes of padding
// behind every second slot): the same part of 8 consecutive lanes then lies in 8 distinct 16-byte bank groups (slot start / 16
// mod 8 = 0 4 1 5 2 6 3 7), the quad-wise writes stay conflict-free as well, and the four parts sit at IMMEDIATE offsets of
// one per-lane base address -- no per-trip swizzle arithmetic (MODE 1 spends 21 integer instructions per trip on the
// addresses of its 8 LDS.128). 4608 instead of 4096 bytes per warp.
template <int MODE> __device__ __forceinline__ constexpr unsigned stage_slot(unsigned slot) { return MODE >= 2 ? 4u * slot + (slot >> 1) : 4u * slot; }   // in 16-byte units
template <int MODE>
__device__ __forceinline__ void ld_staged(const uint4 *warp_stage, unsigned which, unsigned lane, uint32_t w[16]) {
    const uint4 *sl = warp_stage + which * stage_slot<MODE>(32u) + stage_slot<MODE>(lane);
    const unsigned z = MODE >= 2 ? 0u : (lane >> 1) & 3u;
    const uint4 a = sl[0u ^ z], b = sl[1u ^ z], c = sl[2u ^ z], d = sl[3u ^ z];
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    w[8] = c.x; w[9] = c.y; w[10] = c.z; w[11] = c.w; w[12] = d.x; w[13] = d.y; w[14] = d.z; w[15] = d.w;
}

//
// OP 1 turns the same kernel into the whole-read backward search of the query_tags path (FastLocate::count_encoded /
// count, r-index.hpp:541-556; src/query_tags.cpp:87-109): every trip is one LF step = the k / s half of a backward
// extension with the count descriptors (a.desc = the desc_c bank), a lane is done when its read is used up or the range
// became empty, and its record {0, len, range.first, size} (empty range: bwt_start 1, size 0 = the reference's {1,0})
// goes to mem_buf[read] directly: one record per read, already in read order.
// Crowded bucket (pgi_layout.h): sl = the lane's staged directory block, z = its swizzle. Index of the chunk block that
// holds bucket offset rel (> 0): first chunk + sub-bucket number (pgi_layout.h).
__device__ __forceinline__ uint32_t staged_chunk_of(const uint4 *sl, unsigned z, uint32_t rel) {
    return sl[0u ^ z].x + (rel >> sl[2u ^ z].z);                          // w0 + (rel >> w10): first chunk + sub-bucket number
}

template <int MODE, int THREADS, int BPS, int OP = 0>
__global__ void __launch_bounds__(THREADS, BPS) fmd_mem_kernel(const MemArgs a) {
    __shared__ uint32_t s_desc[512];
    __shared__ uint4 s_tab[MODE ? 512 : 1];                              // scan tables of the descriptors (pos_rank)
    constexpr unsigned POS = stage_slot<MODE>(32u), ROUND = stage_slot<MODE>(8u);     // staging geometry in 16-byte units: one position (32 slots), one fetch round (8 slots)
    __shared__ uint4 s_stage[MODE ? (THREADS / 32) * 2 * POS : 1];       // per warp: 2 positions x 32 slots x 4 parts of 16 B (MODE 2: + 16 B of padding per two slots)
    __shared__ uint32_t s_blk[THREADS / 32];                              // per warp: run blocks fetched (work counter, kept out of the registers)
    if (threadIdx.x < THREADS / 32) s_blk[threadIdx.x] = 0;
    for (int i = threadIdx.x; i < 512; i += blockDim.x) {
        s_desc[i] = a.desc[i];
        if (MODE) s_tab[i] = reinterpret_cast<const uint4 *>(a.desc + 512)[i];
    }
    __syncthreads();
    const unsigned lane = threadIdx.x & 31u, FULL = 0xFFFFFFFFu;
    const unsigned lt_mask = (1u << lane) - 1u;
    const uint32_t n_reads = (uint32_t) a.n_reads, min_len = (uint32_t) a.min_len;
    uint4 *const wstage = s_stage + (MODE ? (threadIdx.x >> 5) * 2u * POS : 0u);
    // loop-invariant half of the cooperative fetch (MODE 1): where this lane's part of round 0 / position 0 lands in the staging area
    const uint32_t st_dst = (uint32_t) __cvta_generic_to_shared(wstage) + (stage_slot<MODE>(lane >> 2) + (MODE >= 2 ? (lane & 3u) : ((lane & 3u) ^ ((lane >> 3) & 3u)))) * 16u;
    const uint64_t l2pol = MODE >= 4 ? l2_evict_first_policy() : 0ull;
    MemLane L; L.phase = PH_DONE; L.len = 0; L.x = 0; L.j = 0; L.b.k = L.b.kp = 0; L.b.s = 0;
    uint64_t base = 0, cw = 0, cwi = ~0ull;      // cw: cached 8-byte window of the read
    uint32_t rd = 0, wq_next = 0, wq_end = 0, n_ext = 0, nm = 0, n_seed = 0;
    bool exhausted = false, have = false;
    // 1: the seed lookup of the pending walk was tried and its entry fails the occurrence test: extend step by step;
    // 2: the read holds a byte other than A / C / G / T: no lookups at all (until the next read)
    uint32_t stepwise = 0;
    const uint32_t seed_k = OP == 0 ? a.seed_k : 0u;
    for (;;) {
        // ---- hand out reads to idle lanes
        const bool want = !have && !exhausted;
        const unsigned m = __ballot_sync(FULL, want);
        if (m) {
            if (wq_next >= wq_end) {
                unsigned long long got = 0;
                if (lane == 0) got = atomicAdd(a.cursor, 32ull);
                got = __shfl_sync(FULL, got, 0);
                if (got >= n_reads) { wq_next = wq_end = n_reads; exhausted = true; }
                else { wq_next = (uint32_t) got; wq_end = got + 32 < n_reads ? (uint32_t) got + 32 : n_reads; }
            }
            if (want && !exhausted) {
                const uint32_t id = wq_next + __popc(m & lt_mask);
                if (id < wq_end) {
                    rd = id; have = true; nm = 0;
                    if (OP == 0 && seed_k) stepwise = ((__ldg(a.read_flags + (id >> 5)) >> (id & 31u)) & 1u) ? 2u : 0u;
                    uint64_t o0 = __ldcs(a.read_off + id), o1 = __ldcs(a.read_off + id + 1);
                    if (o1 < o0 || o1 > a.reads_bytes) {                    // offsets that are not ascending or leave the buffer: an empty
                        if (a.bad_off) *a.bad_off = 1ull;                   // read instead of a wild address; the host reports PGI_EINVAL
                        o0 = o1 = 0;
                    }
                    base = o0;
                    if (OP == 0) lane_begin(L, o1 - o0, min_len, a.ix.n);
                    else {
                        L.len = (uint32_t) (o1 - o0 < PGI_MAX_READ_LEN ? o1 - o0 : PGI_MAX_READ_LEN);
                        L.x = 0; L.j = L.len - 1u; L.phase = PH_P1;                      // range = {0, n - 1}; walk i = len .. 1
                        L.b.k = 0; L.b.kp = 0; L.b.s = (int64_t) a.ix.n;
                    }
                    const char *rp = reinterpret_cast<const char *>(a.reads64) + base;      // pull the read towards L2
                    if (!(OP == 0 && seed_k && !stepwise)) for (uint32_t o = 0; o < L.len && o < 512; o += 128) prefetch_l2(rp + o);
                    if (OP == 0 && seed_k && !stepwise) {                   // or its 2-bit codes (a quarter of the bytes): symbols and seed lookups read those
                        const char *cp = reinterpret_cast<const char *>(a.codes32) + (base >> 2);
                        for (uint32_t o = 0; o < (L.len >> 2) + 4u && o < 1024; o += 128) prefetch_l2(cp + o);
                    }
                }
            }
            const uint32_t adv = wq_next + __popc(m);
            wq_next = adv < wq_end ? adv : wq_end;
        }
        // ---- run to the next extension, writing out MEMs on the way
        bool need = false;
        if (have && OP == 1) {
            need = L.phase == PH_P1 && (int32_t) L.j >= 0;
            if (!need) {
                const bool hit = L.phase == PH_P1;
                const uint64_t k = hit ? L.b.k : 1ull, sz = hit ? (uint64_t) L.b.s : 0ull;
                __stcs(a.mem_buf + 2 * (uint64_t) rd, make_uint4(0u, 0u, L.len, 0u));
                __stcs(a.mem_buf + 2 * (uint64_t) rd + 1, make_uint4((uint32_t) k, (uint32_t) (k >> 32), (uint32_t) sz, (uint32_t) (sz >> 32)));
                have = false;
            }
        } else if (have) {
            // lane_settle is straight-line (pgi_core.cuh): one call, at most one MEM to write, and behind an emit the lane
            // extends (min_len >= 1); only min_len = 0 -- a uniform branch -- needs the call repeated
            auto emit = [&]() {
                unsigned long long slot; bool fits;
                if (a.part_cap) {
                    const unsigned part = (blockIdx.x * (THREADS / 32u) + (threadIdx.x >> 5)) & (PGI_SLOT_PARTS - 1u);
                    const unsigned long long k = atomicAdd(a.slot_ctr + 16u * part, 1ull);
                    fits = k < a.part_cap; slot = (unsigned long long) part * a.part_cap + k;
                } else { slot = atomicAdd(a.mem_count, 1ull); fits = slot < a.mem_cap; }
                if (fits) {
                    PGI_CHECK(slot < (a.part_cap ? a.part_cap * PGI_SLOT_PARTS : a.mem_cap)); PGI_CHECK(rd < a.n_reads);
                    __stcs(a.mem_buf + 2 * slot, make_uint4(L.x, 0u, L.j, 0u));
                    __stcs(a.mem_buf + 2 * slot + 1, make_uint4((uint32_t) L.b.k, (uint32_t) (L.b.k >> 32), (uint32_t) (uint64_t) L.b.s, (uint32_t) ((uint64_t) L.b.s >> 32)));
                    __stcs(a.mem_key + slot, ((uint64_t) rd << 32) | nm);
                }
                ++nm;
                lane_emitted(L, a.ix.n);
            };
            int st = lane_settle(L, min_len, a.ix.n);
            if (st == LANE_EMIT) {
                emit(); st = LANE_EXTEND;
                if (min_len == 0) while ((st = lane_settle(L, min_len, a.ix.n)) == LANE_EMIT) emit();
            }
            need = st == LANE_EXTEND;
            if (!need) { a.read_nmems[rd] = nm; have = false; }
        }
        if (__all_sync(FULL, !have && exhausted)) break;
        // ---- one FMD extension
        // forward extension = backward extension of the swapped interval with the complement's descriptor
        // (src/r-index.cpp:758-764): ONE call site, so lanes in different phases share the same instructions
        // The 8-byte window of the read holding the symbol is REQUESTED here and USED only behind the block fetch below, so
        // that its latency overlaps the blocks' instead of preceding it (ncu r01g: a quarter of the stall samples sat on the
        // extraction of the symbol, in front of the cp.async sequence).
        // Seed lookup (pgi_core.cuh "seed table"): a lane whose walk starts from the full interval with seed_k symbols in
        // reach fetches ONE table entry in this trip instead of its run blocks and advances seed_k steps at once. The
        // window of the read is requested here; the index is formed and the entry's cp.async joins the block fetch's
        // group below, so the lookup costs no dependent latency beyond the one the trip pays anyway.
        bool cand = false; uint32_t sc0 = 0, sc1 = 0; uint64_t slo = 0;
        if (OP == 0 && seed_k) {
            cand = need && !stepwise && lane_seed_ready(L, seed_k, a.ix.n);
            // no branch around the two loads: every lane loads, the others from the start of the code stream (one line for the
            // whole warp) -- the divergent form ran at 1.9 active threads in three trips out of four
            slo = cand ? base + L.j + 1u - seed_k : 0ull;                    // first byte of the window
            const uint32_t *p = a.codes32 + (slo >> 4);                      // bit 2 lo of the code stream
            sc0 = __ldg(p); sc1 = __ldg(p + 1);
            need = need && !cand;                                             // no run blocks for a candidate in this trip
        }
        // The symbol comes from the lane's cached 8-byte window of the read. A read of nothing but A / C / G / T (no flag from
        // pack_reads_kernel) is walked through its 2-BIT CODES instead of its bytes: one window then holds 32 symbols instead
        // of 8. The L2 turns over every 25 us under this kernel (5 TB/s through 126 MB), so every new window of a read is a
        // DRAM access of its own: 19 per pass over a 150-base read as bytes, 5 as codes (reads + codes were 11 % of the
        // kernel's DRAM traffic, profiles/r02r).
        uint32_t sym = 0;
        const bool fwd = OP == 0 && L.phase == PH_P2;
        const bool has_sym = need && L.j < L.len;
        const bool by_code = OP == 0 && seed_k != 0 && !(stepwise & 2u);
        {
            const uint64_t o = base + L.j;
            PGI_CHECK(!has_sym || o < a.reads_bytes);
            const uint64_t wi = by_code ? (o >> 5) | (1ull << 63) : o >> 3;           // the tag keeps the two kinds of window apart
            const bool miss = has_sym && wi != cwi;
            const uint64_t *src = by_code ? reinterpret_cast<const uint64_t *>(a.codes32) + (o >> 5) : a.reads64 + (o >> 3);
            if (miss) cw = __ldcs(src);
            cwi = miss ? wi : cwi;
            if (miss && L.len > 512) prefetch_l2(src + (fwd ? 16 : -16));             // long reads: stay ahead in the walking direction
        }
        auto symbol_of = [&](uint64_t win) -> uint32_t {
            const uint32_t o = (uint32_t) (base + L.j);
            const uint32_t as_code = (0x47544341u >> (8u * ((uint32_t) (win >> (2u * (o & 31u))) & 3u))) & 0xFFu;   // "ACTG"[code] (pgi_core.cuh seed_symbol)
            const uint32_t as_byte = (uint32_t) (win >> ((o & 7u) * 8u)) & 0xFFu;
            return by_code ? as_code : as_byte;
        };
        if (MODE == 0 && has_sym) sym = symbol_of(cw);
        if (MODE == 0 && OP == 0 && cand) {
            bool ok = true;
            {
                const uint4 e = __ldg(a.seed + seed_index_packed(sc0, sc1, base + L.j + 1u - seed_k, seed_k));
                SeedWord q; q.x = e.x; q.y = e.y; q.z = e.z; q.w = e.w;
                const BiInt t = seed_unpack(q);
                ok = !ext_failed(t, a.min_occ);
                if (ok) { lane_seed_apply(L, t, seed_k, min_len); n_ext += seed_k; ++n_seed; }
            }
            stepwise = ok ? 0u : 1u;
        }
        if (MODE == 0) {
            if (need) {
                stepwise &= 2u;
                BiInt t; t.k = fwd ? L.b.kp : L.b.k; t.kp = fwd ? L.b.k : L.b.kp; t.s = L.b.s;
                t = extend_desc(a.ix, t, s_desc[(fwd ? 256u : 0u) + sym]);
                BiInt u; u.k = fwd ? t.kp : t.k; u.kp = fwd ? t.k : t.kp; u.s = t.s;
                ++n_ext;
                if (OP == 0) lane_after_ext(L, u, min_len, a.min_occ, a.ix.n);
                else if (ext_failed(u, 0)) L.phase = PH_DONE;                            // LF returned {1, 0}: stays empty
                else { L.b = u; --L.j; }
            }
        } else {
            // (1) the block of a position is its bucket number: no directory access (pgi_layout.h)
            uint32_t b1 = PGI_NOBLK, b2 = PGI_NOBLK;       // b2 stays PGI_NOBLK when k and k+s share a bucket
            if (need) {
                uint64_t p1 = fwd ? L.b.kp : L.b.k, p2 = p1 + (uint64_t) L.b.s;
                if (p2 > a.ix.n) p2 = a.ix.n;
                if (p1 > a.ix.n) p1 = a.ix.n;
                b1 = (uint32_t) (p1 >> a.ix.bshift);
                const uint32_t bb = (uint32_t) (p2 >> a.ix.bshift);
                if (bb != b1) b2 = bb;
            }
            // (2) the warp fetches the blocks together: quad q of round r serves slot 8 r + q
            const unsigned part = lane & 3u;
            const char *const blk_bytes = reinterpret_cast<const char *>(a.ix.blocks);
            // slot 8 r + q, part p lands at byte 512 r + 64 q + 16 (p ^ swizzle(q)): the swizzle does not depend on r.
            // Block indices travel as 32-bit block numbers (n_blocks < 2^32 - 1 is checked at load: 256 GB of run blocks,
            // more than the HBM holds); the receiving lane forms the address of its 16-byte part with one 64-bit multiply-add.
            const char *const part_bytes = blk_bytes + part * 16u;
            auto coop_issue = [&](uint32_t i1, uint32_t i2) {
#pragma unroll
                for (unsigned r = 0; r < 4; ++r) {
                    const unsigned src = 8u * r + (lane >> 2);
                    const uint32_t x1 = __shfl_sync(FULL, i1, src), x2 = __shfl_sync(FULL, i2, src);
                    if (MODE >= 4) {
                        cp_async16_blk_ef(st_dst + 16u * ROUND * r, part_bytes, x1, l2pol);
                        cp_async16_blk_ef(st_dst + 16u * POS + 16u * ROUND * r, part_bytes, x2, l2pol);
                    } else if (MODE >= 2) {
                        cp_async16_blk(st_dst + 16u * ROUND * r, part_bytes, x1);
                        cp_async16_blk(st_dst + 16u * POS + 16u * ROUND * r, part_bytes, x2);
                    } else {
                        cp_async16_if(st_dst + 512u * r, part_bytes + (uint64_t) x1 * 64u, x1);
                        cp_async16_if(st_dst + 2048u + 512u * r, part_bytes + (uint64_t) x2 * 64u, x2);
                    }
                }
            };
            auto coop_fetch = [&](uint32_t i1, uint32_t i2) { coop_issue(i1, i2); cp_async_wait_all(); __syncwarp(); };
            PGI_CHECK(b1 == PGI_NOBLK || b1 < a.ix.n_buckets_dbg); PGI_CHECK(b2 == PGI_NOBLK || b2 < a.ix.n_buckets_dbg);
            coop_issue(b1, b2);
            {
                const unsigned nb = __popc(__ballot_sync(FULL, b1 != PGI_NOBLK)) + __popc(__ballot_sync(FULL, b2 != PGI_NOBLK));
                if (lane == 0) s_blk[threadIdx.x >> 5] += nb;
            }
            if (MODE && OP == 0 && seed_k) {
                // every lane forms an index (a non-candidate from the start of the code stream) and the copy is predicated:
                // no divergent block. Destination: the first 16 bytes of the lane's OWN staging slot -- a candidate fetches
                // no run block in this trip (b1 = b2 = PGI_NOBLK, no chunk either), so nobody else writes there
                const uint32_t sidx = seed_index_packed(sc0, sc1, slo, seed_k);
                PGI_CHECK(sidx < (1u << (2u * seed_k))); PGI_CHECK(!cand || (base + L.j + 1u >= seed_k && base + L.j < a.reads_bytes));
                const uint32_t sdst = (uint32_t) __cvta_generic_to_shared(wstage + stage_slot<MODE>(lane));
                if (MODE >= 4) cp_async16_if_ef(sdst, a.seed + sidx, cand ? 0u : PGI_NOBLK, l2pol);
                else cp_async16_if(sdst, a.seed + sidx, cand ? 0u : PGI_NOBLK);
            }
            cp_async_wait_all();
            __syncwarp();
            // (2b) crowded buckets: a staged block that is a DIRECTORY block (pgi_layout.h) is replaced by the chunk block
            // holding the position, through one more cooperative fetch that the whole warp takes part in, so that the scan
            // below stays ONE converged code path (a per-lane slow path here ran at 1.8 active lanes and cost half as many
            // warp instructions as the scan itself: every fresh interval starts at position 0 of the crowded bucket 0).
            // Offset 0 needs no chunk: the directory block's own header holds the counts before the bucket.
            // (Taking this second fetch OUT of the trip -- the lane remembers its chunk blocks, skips the scan and fetches them
            // as its blocks in the next trip -- was measured: 8.67 ms against 8.47 ms, profiles/r02z: the kernel is bound by
            // DRAM accesses, not by the dependent round trip, and the deferred lanes fetch their other block twice. So was
            // fetching the chunk blocks lane-privately -- four predicated 16-byte copies per lane off one address instead of the
            // cooperative round: 32 fewer instructions, 7.77 ms against 7.70 ms, profiles/r03h.)
            unsigned which2 = b2 != PGI_NOBLK ? 1u : 0u;
            {
                uint32_t c1 = PGI_NOBLK, c2 = PGI_NOBLK;
                if (need) {
                    uint64_t p1 = fwd ? L.b.kp : L.b.k, p2 = p1 + (uint64_t) L.b.s;
                    if (p2 > a.ix.n) p2 = a.ix.n;
                    if (p1 > a.ix.n) p1 = a.ix.n;
                    const uint32_t mask = (1u << a.ix.bshift) - 1u, rel1 = (uint32_t) p1 & mask, rel2 = (uint32_t) p2 & mask;
                    const unsigned z = MODE >= 2 ? 0u : (lane >> 1) & 3u;
                    const uint4 *sl1 = wstage + stage_slot<MODE>(lane), *sl2 = wstage + which2 * POS + stage_slot<MODE>(lane);
                    if (rel1 && (sl1[2u ^ z].y & 0xFu) == PGI_CODE_DIRBLOCK) c1 = staged_chunk_of(sl1, z, rel1);
                    if (rel2 && (sl2[2u ^ z].y & 0xFu) == PGI_CODE_DIRBLOCK) c2 = staged_chunk_of(sl2, z, rel2);
                }
                if (__any_sync(FULL, (c1 & c2) != PGI_NOBLK)) {
                    __syncwarp();                             // every lane has read its directory words
                    PGI_CHECK(c1 == PGI_NOBLK || (c1 >= a.ix.n_buckets_dbg && c1 < a.ix.n_blocks)); PGI_CHECK(c2 == PGI_NOBLK || (c2 >= a.ix.n_buckets_dbg && c2 < a.ix.n_blocks));
                    const unsigned nc = __popc(__ballot_sync(FULL, c1 != PGI_NOBLK)) + __popc(__ballot_sync(FULL, c2 != PGI_NOBLK));
                    if (lane == 0) s_blk[threadIdx.x >> 5] += nc;
                    coop_fetch(c1, c2);
                    if (c2 != PGI_NOBLK) which2 = 1u;
                }
            }
            asm volatile("" : "+l"(cw));                      // pins the use of the read window behind the wait
            if (has_sym) sym = symbol_of(cw);
            // (3) every lane with work scans its own blocks out of shared memory
            if (need) {
                BiInt t; t.k = fwd ? L.b.kp : L.b.k; t.kp = fwd ? L.b.k : L.b.kp; t.s = L.b.s;
                uint64_t p1 = t.k, p2 = p1 + (uint64_t) t.s;
                if (p2 > a.ix.n) p2 = a.ix.n;
                if (p1 > a.ix.n) p1 = a.ix.n;
                const uint32_t di = (fwd ? 256u : 0u) + sym, d = s_desc[di], tbl = d & 0xFFFFFFu;
                const uint4 tq = s_tab[di];
                ScanTab T; T.x = tq.x; T.y = tq.y; T.z = tq.z; T.w = tq.w;
                const uint32_t mask = (1u << a.ix.bshift) - 1u;
                uint32_t w[16];
                ld_staged<MODE>(wstage, 0u, lane, w);
                const PosRank r1 = pos_rank(w, (uint32_t) p1 & mask, tbl, T);
                ld_staged<MODE>(wstage, which2, lane, w);
                const PosRank r2 = pos_rank(w, (uint32_t) p2 & mask, tbl, T);
                t = extend_finish(a.ix, t, d, r1, r2, p1, p2);
                BiInt u; u.k = fwd ? t.kp : t.k; u.kp = fwd ? t.k : t.kp; u.s = t.s;
                ++n_ext;
                if (OP == 0) lane_after_ext(L, u, min_len, a.min_occ, a.ix.n);
                else if (ext_failed(u, 0)) L.phase = PH_DONE;                            // LF returned {1, 0}: stays empty
                else { L.b = u; --L.j; }
                stepwise &= 2u;                               // whatever walk comes next may use the table again
            }
            if (OP == 0 && seed_k) {                          // uniform; the candidates take their entry by selects
                const uint4 e = wstage[stage_slot<MODE>(lane)];
                SeedWord q; q.x = e.x; q.y = e.y; q.z = e.z; q.w = e.w;
                const BiInt t = seed_unpack(q);
                const bool ok = cand && !ext_failed(t, a.min_occ);
                L.b.k = ok ? t.k : L.b.k; L.b.kp = ok ? t.kp : L.b.kp; L.b.s = ok ? t.s : L.b.s;   // lane_seed_apply
                L.j = ok ? L.j - seed_k : L.j;
                n_ext += ok ? seed_k : 0u; n_seed += ok ? 1u : 0u;
                stepwise = cand ? (ok ? 0u : 1u) : stepwise;  // too few occurrences: the walk goes step by step
            }
            __syncwarp();                                     // the staging area is rewritten by the next trip
        }
    }
    unsigned long long tot = n_ext;
    for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(FULL, tot, o);
    if (lane == 0 && tot) atomicAdd(a.ext_count, tot);
    if (MODE && lane == 0 && a.blk_count && s_blk[threadIdx.x >> 5]) atomicAdd(a.blk_count, (unsigned long long) s_blk[threadIdx.x >> 5]);
    if (OP == 0 && seed_k) {
        unsigned tots = n_seed;
        for (int o = 16; o; o >>= 1) tots += __shfl_xor_sync(FULL, tots, o);
        if (lane == 0 && tots) atomicAdd(a.seed_count, (unsigned long long) tots);
    }
}

// K0. Streaming pre-pass for the seed lookups (pgi_core.cuh "seed table"): one thread per aligned 8-byte word of the reads
// buffer writes the word's eight 2-bit codes; a byte other than A / C / G / T (rare: N, lower case) flags the read that
// holds it (binary search over read_off, only then). codes16 has n_words + 4 units (the lookups load two 32-bit units).
__global__ void pack_reads_kernel(const uint64_t *reads64, uint64_t n_words, uint64_t n_bytes, const uint64_t *read_off, uint64_t n_reads,
                                  uint16_t *codes16, uint32_t *read_flags) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i >= n_words) return;
    uint32_t bad;
    codes16[i] = (uint16_t) seed_pack8(__ldcs(reads64 + i), bad);
    while (bad) {
        const uint32_t bi = __ffs(bad) - 1u; bad &= bad - 1u;
        const uint64_t pos = 8u * i + bi;
        if (pos >= n_bytes) break;                                       // padding behind the last read
        uint64_t lo = 0, hi = n_reads;                                   // last r with read_off[r] <= pos
        while (hi - lo > 1) { const uint64_t mid = (lo + hi) >> 1; if (__ldg(read_off + mid) <= pos) lo = mid; else hi = mid; }
        atomicOr(read_flags + (lo >> 5), 1u << (lo & 31u));
    }
}

// K4. Appended MEMs -> read order: dst = mem_off[read] + index within the read. The number of appended MEMs is read from
// device memory (state[0]; the host does not know it yet: the whole step is stream-ordered, pgi_capi.cu batch_run).
__global__ void mem_scatter_kernel(const uint4 *mem_buf, const uint64_t *mem_key, const uint64_t *mem_off,
                                   uint64_t n_slots, uint4 *mems, const unsigned long long *slot_ctr, uint64_t part_cap,
                                   const uint64_t *n_appended, uint64_t mems_cap) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i >= n_slots) return;                                           // partitioned: n_slots = PGI_SLOT_PARTS * part_cap slots
    if (part_cap) { if (i % part_cap >= slot_ctr[16u * (i / part_cap)]) return; }   // behind the filled prefix of its partition
    else if (i >= *n_appended) return;
    const uint64_t key = mem_key[i];
    const uint64_t dst = mem_off[key >> 32] + (key & 0xFFFFFFFFull);
    if (dst >= mems_cap) return;                                        // only after an overflow (the step is run again)
    mems[2 * dst] = mem_buf[2 * i]; mems[2 * dst + 1] = mem_buf[2 * i + 1];
}

// Batched backward / forward extension (FastLocate::backward_extend_encoded / forward_extend_encoded).
__global__ void fmd_extend_kernel(RiDev ix, const uint32_t *desc, uint64_t n, const uint64_t *in3, const uint8_t *sym,
                                  const uint8_t *dir, uint64_t *out3) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i >= n) return;
    BiInt b; b.k = in3[3 * i]; b.kp = in3[3 * i + 1]; b.s = (int64_t) in3[3 * i + 2];
    if (dir[i]) {
        BiInt t; t.k = b.kp; t.kp = b.k; t.s = b.s;
        t = extend_desc(ix, t, desc[256 + sym[i]]);
        b.k = t.kp; b.kp = t.k; b.s = t.s;
    } else b = extend_desc(ix, b, desc[sym[i]]);
    out3[3 * i] = b.k; out3[3 * i + 1] = b.kp; out3[3 * i + 2] = (uint64_t) b.s;
}

// Batched rank vectors R(pos)[0..8) (FastLocate::rank_at_cached_encoded); slot = RiConsts::slot.
struct SlotTable { uint32_t s[8]; uint32_t sigma; };
__global__ void rank_kernel(RiDev ix, SlotTable st, uint64_t n, const uint64_t *pos, uint64_t *out8) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t p = pos[i]; if (p > ix.n) p = ix.n;
    uint64_t t1[6];
    rank_all(ix, p, t1);
    const uint64_t q = ix.has_quirk ? quirk_at(ix, p) : 0;
    for (uint32_t j = 0; j < 8; ++j) {
        uint64_t v = 0;
        if (j < st.sigma) {
            const uint32_t s = st.s[j];
            if (s == PGI_SRC_QUIRK) v = q;
            else if (s < 6) { v = 0; for (int c = 0; c < 6; ++c) if (s == (uint32_t) c) v = t1[c]; }
        }
        out8[8 * i + j] = v;
    }
}

} // namespace pgi
