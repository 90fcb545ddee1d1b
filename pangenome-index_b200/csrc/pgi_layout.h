// pgi_layout.h — constants and plain structs shared by the host flattener and the device code.
// See DESIGN.md "Data layout in HBM" for the rationale.
#pragma once
#include <stdint.h>

// One run block = 64 bytes = 8 little-endian u64 words (two 32-byte DRAM sectors, one aligned 64 B burst):
//   w0 = start(40) | A[0..24)           start = first BWT position covered by the block
//   w1 = A[24..40) | C(40) | spare(8)   A,C,G,T = symbol counts before the block (40 bits)
//   w2 = G(40) | T[0..24)
//   w3 = T[24..40) | NL(32) | N[0..16)  NL = endmarker count, N = 'N' count (32 bits)
//   w4 = N[16..32) | e0 e1 e2           e_i = 16-bit run entries: code(3) | length(13), length 0 = unused
//   w5 = e3..e6, w6 = e7..e10, w7 = e11..e14
// code = index into nuc = {'\n','A','C','G','N','T'} (include/pangenome_index/utils.hpp:11).
#define PGI_BLOCK_WORDS 8
#define PGI_RUNS_PER_BLOCK 15
#define PGI_MAX_RUN_ENTRY 8191ull

// Sources of one rank-vector slot R(pos)[i]: true rank of nuc code 0..5,
// the legacy "column-0 cumulative at the reference block start" quirk, or constant zero.
#define PGI_SRC_QUIRK 6u
#define PGI_SRC_ZERO 7u

// Per-byte extension descriptor (desc_b for backward, desc_f = desc_b[complement] for forward):
//   bits 0..17  w[c], c = 0..5 : how many times (true rank delta of code c) is added to k'
//   bits 18..20 wq             : how many times the quirk delta is added to k'
//   bits 21..23 main source    : where R[sym_map[a]] comes from (0..5, PGI_SRC_QUIRK, PGI_SRC_ZERO)
//   bits 24..26 sym_map[a]     : index into C
typedef struct {
    uint64_t n, sigma, n_blocks, n_quirk, mem_tail;
    uint32_t dir_shift, has_quirk;
    uint32_t slot[8];          // source of R(pos)[i], i < sigma (0..5 true rank of that nuc code, PGI_SRC_QUIRK, PGI_SRC_ZERO)
    uint64_t C[8];
    uint8_t sym_map[256];
    uint32_t desc_b[256], desc_f[256];
} RiConsts;

// Sorted set S of positions in [0,size): rank1(i) = |{p in S : p < i}|.
// dir[b] = |{p in S : p < b << shift}| (n_buckets + 1 entries); low[j] = j-th element & (2^shift - 1),
// stored in low_bytes bytes. Replaces sdsl::sd_vector rank_1 (src/tag_arrays.cpp:857-858).
typedef struct {
    uint64_t size, m, n_buckets;
    uint32_t shift, low_bytes;
} SparseRankConsts;
