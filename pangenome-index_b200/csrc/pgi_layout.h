// pgi_layout.h — constants and plain structs shared by the host flattener and the device code.
// See DESIGN.md "Data layout in HBM" for the rationale.
#pragma once
#include <stdint.h>

// The BWT is cut into BUCKETS of 2^bshift consecutive positions and bucket b is the 64-byte block number b: the block
// of a position is pos >> bshift -- no directory, no predecessor search, ONE random 64-byte access per rank query
// (the reference's blocks_start_pos.predecessor + block decode, src/r-index.cpp:594,621, costs 3-4 dependent lines).
// Runs are clipped at bucket boundaries (a run that crosses one appears in both buckets).
//
// One block = 64 bytes = 16 little-endian u32 words (two 32-byte DRAM sectors, one aligned 64 B burst):
//   w0      normal block: 0;  directory block (below): index of its first chunk block
//   w1..w8  the symbol counts before the first position of the block's runs (A, C, G, T: 40 bits; NL, N: 32 bits), cut into
//           LIMBS and stored in code order ('\n' A C G N T = codes 0..5), so that the weighted sum over the six counts is a
//           handful of byte dot products against the same 8-byte weight tables the run scan uses (pos_rank):
//   w1..w3  bits 0..15 of  NL | A << 16,   C | G << 16,   N | T << 16
//   w4..w6  bits 16..31 of the same pairs
//   w7      bits 32..39:   0 | A << 8 | C << 16 | G << 24
//   w8                     0 | T << 8 | codes 8..11 << 16         (one nibble per run)
//   w9      codes 0..7                                          (one nibble per run, run 0 in the low nibble)
//   w10..w15 E_0..E_11, two per word (even run in the low half)
// Run i of the block has symbol code c_i = index into nuc = {'\n','A','C','G','N','T'}
// (include/pangenome_index/utils.hpp:11) and END OFFSET E_i = offset inside the BUCKET of the first position behind
// run i (16 bits). Unused runs have code 7 and repeat the last end offset. Storing ends instead of lengths makes the
// scan data-parallel: the positions of run i that lie before bucket offset rel are min(rel, E_i) - min(rel, E_{i-1}),
// two runs per 16x2 SIMD min, and the weighted sums are byte dot products (dp2a) against 8-byte weight tables looked
// up with the code nibbles as PRMT selectors (pos_rank in pgi_core.cuh).
//
// bshift is chosen per index: the largest value <= 12 for which at most 1/PGI_CROWDED_DIV of the buckets hold more than 12
// clipped runs (PGI_BSHIFT=<n> overrides it for tuning / tests). Larger buckets mean fewer block fetches: the two
// positions k and k + s of an extension share their bucket more often (on a pangenome of H haplotypes s stays near 2 H for
// most of a read), and the B200's DRAM delivers 128 bytes per access whatever is asked for (ncu: 114 B of DRAM reads per
// 64-byte block fetched, profiles/r02h), so every fetch saved is 128 bytes of the bound resource. A bucket with MORE than
// 12 runs becomes a DIRECTORY block:
//   w9 low nibble = 0xF (no run has that code), w0 = index of its first chunk block (chunks are stored behind the
//   n_buckets primary blocks), w10 = sub_shift: the bucket is cut into 2^(bshift - sub_shift) SUB-BUCKETS of 2^sub_shift
//   positions, one chunk block each, sub_shift being the largest value for which every sub-bucket holds at most 11 clipped
//   runs; its header (w1..w8 low half) holds the counts before the bucket, which is the whole answer for bucket offset 0.
// The chunk of bucket offset rel is chunk number rel >> sub_shift: no search. A chunk block has the normal format with
// the counts before ITS first position in the header; chunks 1.. start with a pad entry (code 7, E = the chunk's start
// offset) so that the same scan applies (end offsets stay relative to the BUCKET). A position in such a bucket costs ONE
// more dependent 64-byte access: dense stretches are rare by the choice of bshift (but not negligible: the first n_seq
// BWT rows are runs of length 1 next to C['A'], where every search for an 'A' starts).
#define PGI_BLOCK_U32 16
// Header words of a block from its six counts cnt[c], c = nuc code ('\n' A C G N T). Leaves the code half of w8 alone.
#ifdef __CUDACC__
__host__ __device__
#endif
static inline void pgi_put_header(uint32_t *w, const uint64_t cnt[6]) {
    w[1] = (uint32_t) (cnt[0] & 0xFFFFu) | (uint32_t) (cnt[1] & 0xFFFFu) << 16;
    w[2] = (uint32_t) (cnt[2] & 0xFFFFu) | (uint32_t) (cnt[3] & 0xFFFFu) << 16;
    w[3] = (uint32_t) (cnt[4] & 0xFFFFu) | (uint32_t) (cnt[5] & 0xFFFFu) << 16;
    w[4] = (uint32_t) ((cnt[0] >> 16) & 0xFFFFu) | (uint32_t) ((cnt[1] >> 16) & 0xFFFFu) << 16;
    w[5] = (uint32_t) ((cnt[2] >> 16) & 0xFFFFu) | (uint32_t) ((cnt[3] >> 16) & 0xFFFFu) << 16;
    w[6] = (uint32_t) ((cnt[4] >> 16) & 0xFFFFu) | (uint32_t) ((cnt[5] >> 16) & 0xFFFFu) << 16;
    w[7] = (uint32_t) ((cnt[1] >> 32) & 0xFFu) << 8 | (uint32_t) ((cnt[2] >> 32) & 0xFFu) << 16 | (uint32_t) ((cnt[3] >> 32) & 0xFFu) << 24;
    w[8] = (w[8] & 0xFFFF0000u) | (uint32_t) ((cnt[5] >> 32) & 0xFFu) << 8;
}
#define PGI_BLOCK_WORDS 8            /* the same block counted in u64 words (host vectors are u64) */
#define PGI_RUNS_PER_BLOCK 12
#define PGI_CODE_UNUSED 7u
#define PGI_CODE_DIRBLOCK 0xFu
#define PGI_MAX_BSHIFT 12            /* end offsets are 16 bits; 4096 positions per 64-byte block is already more than 12 runs ever cover */
#define PGI_MIN_BSHIFT 4
#define PGI_CHUNK_RUNS 11            /* runs per chunk block (+ the pad entry) */
#define PGI_CROWDED_DIV 128          /* at most one bucket in this many may be crowded (bshift rule; config 4: shift 9 / 10 / 11 = 9.79 / 9.71 / 9.68 ms, 12: 12.7 ms, profiles/r02i) */

// Sources of one rank-vector slot R(pos)[i]: true rank of nuc code 0..5,
// the legacy "column-0 cumulative at the reference block start" quirk, or constant zero.
#define PGI_SRC_QUIRK 6u
#define PGI_SRC_ZERO 7u

// Per-byte extension descriptor (desc_b for backward, desc_f = desc_b[complement] for forward), 32 bits:
//   bits 0..23  six 4-bit fields, field c (code c): bits 0..2 = w[c], how many times the true rank delta of code c
//               is added to k'; bit 3 = "R[sym_map[a]] is the true rank of code c" (at most one field has it)
//   bits 24..26 wq: how many times the quirk delta is added to k'
//   bit 27      R[sym_map[a]] comes from the quirk slot;  bit 28: it is constant zero
//   bits 29..31 sym_map[a]: index into C
#define PGI_DESC_WORDS (512 + 512 * 4)  /* device copy: desc_b[256], desc_f[256], then a 4-word scan table per descriptor */
#define PGI_DESC_MAIN_QUIRK (1u << 27)
#define PGI_DESC_MAIN_ZERO (1u << 28)
typedef struct {
    uint64_t n, sigma, n_blocks, n_quirk, mem_tail;       // n_blocks: primary (bucket) blocks + chunk blocks
    uint32_t bshift, has_quirk; uint64_t n_buckets, n_dir_blocks;
    uint32_t slot[8];          // source of R(pos)[i], i < sigma (0..5 true rank of that nuc code, PGI_SRC_QUIRK, PGI_SRC_ZERO)
    uint64_t C[8];
    uint8_t sym_map[256];
    uint32_t desc_b[256], desc_f[256];
    // Whole-read backward search (FastLocate::count_encoded / count -> LF_encoded / LF -> rankAt_encoded / rankAt,
    // r-index.hpp:541-556, r-index.cpp:558-590,650-711): one LF step is the k / s half of a backward extension whose
    // "main" rank is the TRUE rank of the byte itself (no rank-vector slot aliasing), so it reuses the same descriptor
    // format with all weights zero. count_ok = 0: encoded file without N, on which the reference reads a run header as
    // a sixth cumulative (r-index.cpp:577) -- not reproduced, the count entry points refuse such an index.
    uint32_t desc_c[256];
    uint32_t count_ok, locate_ok;   // locate_ok: the file carries samples and its run lookup is well defined in the reference
    uint64_t max_length;            // Header::max_length: SA value = seq_id * max_length + seq_offset (r-index.hpp:424-431)
} RiConsts;

// Sorted set S of positions in [0,size): rank1(i) = |{p in S : p < i}|.
// dir[b] = |{p in S : p < b << shift}| (n_buckets + 1 entries); low[j] = j-th element & (2^shift - 1),
// stored in low_bytes bytes. Replaces sdsl::sd_vector rank_1 (src/tag_arrays.cpp:857-858).
typedef struct {
    uint64_t size, m, n_buckets;
    uint32_t shift, low_bytes;
} SparseRankConsts;
