// pgi_layout.h — constants and plain structs shared by the host flattener and the device code.
// See DESIGN.md "Data layout in HBM" for the rationale.
#pragma once
#include <stdint.h>

// One run block = 64 bytes = 16 little-endian u32 words (two 32-byte DRAM sectors, one aligned 64 B burst):
//   w0      start, low 32 bits          start = first BWT position covered by the block (40 bits)
//   w1..w4  A, C, G, T, low 32 bits     symbol counts before the block (40 bits each)
//   w5, w6  NL, N                       endmarker / 'N' counts before the block (32 bits each)
//   w7      start_hi | A_hi << 8 | C_hi << 16 | G_hi << 24      (bits 32..39 of the 40-bit fields)
//   w8      T_hi | spare << 8 | codes 8..11 << 16               (one nibble per run)
//   w9      codes 0..7                                          (one nibble per run, run 0 in the low nibble)
//   w10..w15 E_0..E_11, two per word (even run in the low half)
// Run i of the block has symbol code c_i = index into nuc = {'\n','A','C','G','N','T'}
// (include/pangenome_index/utils.hpp:11) and END OFFSET E_i = number of block positions up to and including run i
// (16 bits, run lengths 1..4095: 12 * 4095 < 2^16). Unused runs have code 7 and repeat the last end offset, so E_11 is
// always the length of the block. Storing ends instead of lengths makes the scan data-parallel: the positions of run i
// that lie before block offset rel are min(rel, E_i) - min(rel, E_{i-1}), two runs per 16x2 SIMD min, and the weighted
// sums are byte dot products (dp2a) against 8-byte weight tables looked up with the code nibbles as PRMT selectors
// (pos_rank in pgi_core.cuh).
#define PGI_BLOCK_U32 16
#define PGI_BLOCK_WORDS 8            /* the same block counted in u64 words (host vectors are u64) */
#define PGI_RUNS_PER_BLOCK 12
#define PGI_MAX_RUN_ENTRY 4095ull
#define PGI_CODE_UNUSED 7u

// Block directory: two levels, both read with independent loads (no dependent chain between them).
//   sb[pos >> sb_shift]      (u32) index of the block containing the first position of the superbucket (2^19..2^30
//                            positions: the table is a handful of cache lines, so 32 lanes' loads share wavefronts)
//   dir[pos >> dir_shift]    (u32) bits 0..15  block index at the bucket start, relative to its superbucket's sb entry
//                                  bits 16..23 offset inside the bucket of the 1st block boundary (0: none)
//                                  bits 24..31 offset of the 2nd block boundary (0: none)
// dir_shift <= 8, chosen so that a bucket is about one block long: 4 bytes per ~block instead of the 8 bytes per half
// block of a one-level table, so the directory of a 2 G-position index is 31 MB and lives in L2 while the blocks stream
// from HBM. A bucket with more than two boundaries makes the guess land short; the scan notices (positions left over
// after the last run) and the rare slow path searches the block starts.
// Dense buckets: a bucket with MORE than two block boundaries inside (short runs: e.g. the first n_seq BWT rows, which
// hold the last symbol of every sequence, next to C['A'] = n_seq where every search for an 'A' starts) gets the
// reserved delta 0xFFFF and, in bits 16..31, the number of a 32-byte record in the dense table:
//   word 0      (u32) absolute index of the block containing the first position of the bucket
//   bytes 4..31 offsets inside the bucket of every block boundary, ascending (0: none)
// so that the guess is exact there too (a bucket holds at most 256 / 12 boundaries). Only when more than 65535
// buckets are dense do the rest keep the two-boundary form and fall back to the search.
#define PGI_DENSE_U32 8
#define PGI_DENSE_BOUNDS 28
#define PGI_SB_MIN_SHIFT 19           /* 2^19 / 12 < 2^16: the 16-bit delta always fits */
#define PGI_SB_MAX_SHIFT 30           /* the flattener takes the largest shift whose deltas still fit: the sb table is then a few cache lines */
#define PGI_DIR_MAX_SHIFT 8

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
    uint64_t n, sigma, n_blocks, n_quirk, mem_tail;
    uint32_t dir_shift, has_quirk, sb_shift, pad_;
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
    uint32_t count_ok, pad2_;
} RiConsts;

// Sorted set S of positions in [0,size): rank1(i) = |{p in S : p < i}|.
// dir[b] = |{p in S : p < b << shift}| (n_buckets + 1 entries); low[j] = j-th element & (2^shift - 1),
// stored in low_bytes bytes. Replaces sdsl::sd_vector rank_1 (src/tag_arrays.cpp:857-858).
typedef struct {
    uint64_t size, m, n_buckets;
    uint32_t shift, low_bytes;
} SparseRankConsts;
