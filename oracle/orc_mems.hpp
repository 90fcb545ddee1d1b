// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_formats.hpp header).
//
// Restatement of include/pangenome_index/algorithm.hpp:644-757 (struct MEM,
// find_mems_function, find_all_mems) and of the per-read loop and stdout grammar of
// src/find_mems.cpp:94-146 (+ tag_arrays.cpp:848-853 / :885-889 for the tag lines).
#pragma once
#include "orc_index.hpp"
#include <string>

namespace orc {

// algorithm.hpp:644-651
struct MEM { uint64_t start, end, bwt_start; int64_t size; };

// Work counters = the roofline denominators of SURVEY.md §8(d).
struct Counters {
    uint64_t extensions = 0;        // E
    uint64_t distinct_blocks = 0;   // D = sum over extensions of |{block(k), block(k+s)}|
    uint64_t mems = 0;              // M
    uint64_t tag_runs = 0;          // sum of number_of_runs
    uint64_t tag_skipped = 0;       // sum of skipped items
    uint64_t tag_item_sectors = 0;  // sum ceil(8*(skip+runs)/32)
    void add(const Counters &o) {
        extensions += o.extensions; distinct_blocks += o.distinct_blocks; mems += o.mems;
        tag_runs += o.tag_runs; tag_skipped += o.tag_skipped; tag_item_sectors += o.tag_item_sectors;
    }
};

// pattern[j] with std::string semantics: pattern[len] == '\0' (algorithm.hpp:722-723)
static inline uint64_t pat_at(const char *pattern, uint64_t len, uint64_t j) {
    return j < len ? (uint64_t) (unsigned char) pattern[j] : 0;
}

// algorithm.hpp:653-736
static inline uint64_t find_mems_function(const char *pattern, uint64_t len, uint64_t min_len, uint64_t min_occ, uint64_t x,
                                          const FastLocate &fmd_index, std::vector<MEM> &output, Counters *ctr) {
    uint64_t j;
    if (len - x < min_len) return len;
    uint64_t d = 0;
    bi_interval bint{0, 0, (int64_t) fmd_index.bwt_size()};
    for (j = x + min_len - 1; (int64_t) j >= (int64_t) x; --j) {
        bint = fmd_index.backward_extend_encoded(bint, pat_at(pattern, len, j), ctr ? &d : nullptr);
        if (ctr) { ctr->extensions++; ctr->distinct_blocks += d; }
        if ((uint64_t) bint.size < min_occ || bint.size <= 0) return j + 1;
        if (j == 0) break;
    }
    bi_interval bint2 = bint;
    for (j = x + min_len; j < len; ++j) {
        bint = fmd_index.forward_extend_encoded(bint, pat_at(pattern, len, j), ctr ? &d : nullptr);
        if (ctr) { ctr->extensions++; ctr->distinct_blocks += d; }
        if ((uint64_t) bint.size < min_occ || bint.size <= 0) break;
        bint2 = bint;
    }
    uint64_t e = j;
    output.push_back(MEM{x, e, bint2.forward, bint2.size});
    bi_interval back{0, 0, (int64_t) fmd_index.bwt_size()};
    for (j = e; j > x; --j) {
        back = fmd_index.backward_extend_encoded(back, pat_at(pattern, len, j), ctr ? &d : nullptr);
        if (ctr) { ctr->extensions++; ctr->distinct_blocks += d; }
        if ((uint64_t) back.size < min_occ || back.size <= 0) return j + 1;
    }
    return j + 1;
}

// algorithm.hpp:739-757 (the per-read stderr line at :754 is not reproduced)
static inline void find_all_mems(const char *pattern, uint64_t len, uint64_t min_len, uint64_t min_occ,
                                 const FastLocate &fmd_index, std::vector<MEM> &mems, Counters *ctr) {
    uint64_t x = 0;
    while (x < len) x = find_mems_function(pattern, len, min_len, min_occ, x, fmd_index, mems, ctr);
}

// Text block of one read, src/find_mems.cpp:115-138 + tag_arrays.cpp:885-889.
static inline void format_read(std::string &out, uint64_t seq_no, const MEM *mems, uint64_t n_mems,
                               const uint64_t *tag_off, const uint64_t *tags, bool with_tags) {
    out += "Seq: "; out += std::to_string(seq_no); out += '\n';
    for (uint64_t m = 0; m < n_mems; ++m) {
        out += "MEM START: "; out += std::to_string(mems[m].start);
        out += ", MEM END: "; out += std::to_string(mems[m].end);
        out += " BWT START: "; out += std::to_string(mems[m].bwt_start);
        out += " SIZE: "; out += std::to_string(mems[m].size); out += '\n';
        if (with_tags) {
            uint64_t a = tag_off[m], b = tag_off[m + 1];
            out += "Number of unique positions: "; out += std::to_string(b - a); out += '\n';
            for (uint64_t t = a; t < b; ++t) { out += std::to_string(tags[t]); out += ", "; }
            out += '\n';
        }
    }
    out += '\n';
}

} // namespace orc
