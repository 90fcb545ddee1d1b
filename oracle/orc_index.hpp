// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_formats.hpp header).
//
// Line-by-line CPU restatement of the reference's query-side index classes:
//   panindexer::FastLocate  (src/r-index.cpp, include/pangenome_index/r-index.hpp)
//   panindexer::TagArray    (src/tag_arrays.cpp, include/pangenome_index/tag_arrays.hpp)
// Every function cites the reference lines it follows. All quirks listed in
// SURVEY.md §0 (sigma=5 slot aliasing, pattern[len], tag start off-by-one) are
// reproduced literally; behaviour the reference leaves undefined is defined here
// and listed in DESIGN.md ("defined UB").
//
// Parity status: "parity unpinned" by any reference golden output for find_mems
// (the reference ships none and cannot be built here); pinned by the shipped
// binary fixtures (layouts), the FMD properties of tests/test_rindex.cpp:288-486,
// brute-force MEM/occurrence checks on sigma=6 texts and SURVEY.md App. C/G vectors.
#pragma once
#include "orc_formats.hpp"
#include <array>
#include <utility>

namespace orc {

// include/pangenome_index/utils.hpp:9-11
static const unsigned char NENDMARKER = '\n';
static const unsigned char NUC[6] = {'\n', 'A', 'C', 'G', 'N', 'T'};

// r-index.hpp:118-130
struct bi_interval { uint64_t forward, reverse; int64_t size; };

typedef std::array<uint64_t, 8> rankvec;   // the reference's std::vector<size_t>(C.size()); C.size() <= 6 enforced at load

struct FastLocate {
    // Header, r-index.hpp:85-115
    uint32_t tag = 0, version = 0; uint64_t max_length = 0, flags = 0;
    IntVec samples; SdVec last; IntVec last_to_run;
    uint8_t sym_map[256]; std::vector<uint64_t> C;
    SdVec blocks_start_pos; uint64_t sequence_size = 0;
    // legacy blocks, r-index.hpp:134-297
    struct RunBlock { std::vector<uint64_t> cum; std::vector<std::pair<uint64_t, uint64_t>> runs; };
    std::vector<RunBlock> blocks;
    // encoded blocks, r-index.hpp:320-335
    uint64_t encoded_block_size = 0; bool encoded_has_N = false;
    IntVec blocks_encoded_start_bits; std::vector<uint8_t> blocks_encoded_stream;
    uint8_t complement_table[256];

    static const uint32_t TAG = 0x6B3741D8u;

    uint64_t bwt_size() const { return sequence_size; }
    bool is_encoded() const { return !blocks_encoded_start_bits.empty(); }     // r-index.hpp:412
    uint64_t complement(uint64_t a) const { return complement_table[a & 0xFF]; }

    // r-index.cpp:1512-1529
    void initialize_complement_table() {
        for (int i = 0; i < 256; ++i) complement_table[i] = (uint8_t) i;
        complement_table['A'] = 'T'; complement_table['C'] = 'G'; complement_table['G'] = 'C'; complement_table['T'] = 'A';
        complement_table['a'] = 't'; complement_table['c'] = 'g'; complement_table['g'] = 'c'; complement_table['t'] = 'a';
    }

    void load_common(Cursor &c) {
        samples.load(c, 0); last.load(c); last_to_run.load(c, 0);
        IntVec sm; sm.load(c, 8);
        if (sm.size() != 256) throw std::runtime_error("oracle: sym_map must have 256 entries");
        for (int i = 0; i < 256; ++i) sym_map[i] = (uint8_t) sm.get(i);
        IntVec Cv; Cv.load(c, 64);
        C.resize(Cv.size()); for (size_t i = 0; i < C.size(); ++i) C[i] = Cv.get(i);
        if (C.empty() || C.size() > 6) throw std::runtime_error("oracle: alphabet size outside [1,6] is undefined in the reference (nuc has 6 entries)");
        blocks_start_pos.load(c);
        sequence_size = c.u64();
    }

    // r-index.cpp:406-459 (load_encoded, falling back to load :378-404 when flag bit0 is clear)
    void load_encoded(Cursor &c) {
        initialize_complement_table();
        tag = c.u32(); version = c.u32(); max_length = c.u64(); flags = c.u64();
        if (tag != TAG) throw std::runtime_error("FastLocate: Invalid tag");
        if (version != 1) throw std::runtime_error("FastLocate: Expected v1, got v" + std::to_string(version));
        load_common(c);
        if ((flags & 1) == 0) {
            uint64_t nb = c.u64();
            blocks.resize(nb);
            for (uint64_t i = 0; i < nb; ++i) {
                IntVec cum; cum.load(c, 64);
                blocks[i].cum.resize(cum.size());
                for (size_t j = 0; j < blocks[i].cum.size(); ++j) blocks[i].cum[j] = cum.get(j);
                uint64_t nr = c.u64();
                blocks[i].runs.resize(nr);
                for (uint64_t j = 0; j < nr; ++j) { blocks[i].runs[j].first = c.u64(); blocks[i].runs[j].second = c.u64(); }
            }
            return;
        }
        encoded_block_size = c.u64();
        encoded_has_N = c.u8() != 0;
        blocks_encoded_start_bits.load(c, 0);
        uint64_t nbytes = c.u64();
        const uint8_t *src = c.bytes(nbytes);
        blocks_encoded_stream.assign(src, src + nbytes);
    }

    // sd_vector::predecessor as used at r-index.cpp:594,621: (block id, block start).
    // pos >= size resolves to the last block (needed for pos == n on the initial interval;
    // pos > n only arises from quirk arithmetic and is defined the same way).
    inline void locate_block(uint64_t pos, uint64_t &bid, uint64_t &bstart) const {
        uint64_t j = blocks_start_pos.predecessor_index(pos);
        if (j == UINT64_MAX) { bid = 0; bstart = 0; return; }   // no block: empty index
        bid = j; bstart = blocks_start_pos.ones[j];
    }

    // Run_blocks::rankAt, r-index.hpp:180-221 (count of `symbol` among the first `pos` block positions)
    static uint64_t block_rankAt(const RunBlock &b, uint64_t pos, uint64_t symbol) {
        uint64_t rank = 0, current_position = 0; size_t run_num = 0;
        while (run_num < b.runs.size()) {
            if (b.runs[run_num].first == symbol) {
                if (current_position + b.runs[run_num].second > pos) { rank += (pos - current_position); break; }
                else rank += b.runs[run_num].second;
            }
            current_position += b.runs[run_num].second;
            run_num++;
            if (current_position > pos) break;
        }
        return rank;
    }

    // r-index.cpp:593-603
    rankvec rank_at_cached(uint64_t pos) const {
        rankvec r{}; uint64_t bid, bstart; locate_block(pos, bid, bstart);
        if (bid >= blocks.size()) return r;
        const RunBlock &b = blocks[bid];
        for (size_t i = 0; i < C.size(); ++i) {
            uint64_t cum_rank = block_rankAt(b, pos - bstart, NUC[i]);
            uint64_t col = sym_map[NUC[i]];
            r[i] = cum_rank + (col < b.cum.size() ? b.cum[col] : 0);
        }
        return r;
    }

    // EncodedBlock::read_cumulative, r-index.cpp:65-82
    void read_cumulative(size_t &loc, size_t cum_len, uint64_t cum_nuc[6]) const {
        for (int i = 0; i < 6; ++i) cum_nuc[i] = 0;
        for (size_t idx = 0; idx < cum_len; ++idx) {
            uint64_t val = bytecode_read(blocks_encoded_stream.data(), blocks_encoded_stream.size(), loc);
            for (int i = 0; i < 6; ++i) if (sym_map[NUC[i]] == idx) { cum_nuc[i] = val; break; }
        }
        if (!encoded_has_N) cum_nuc[4] = 0;
    }

    // EncodedBlock::ranks_at, r-index.cpp:133-156
    void ranks_at(size_t loc, size_t end_pos, uint64_t rel, size_t cum_len, uint64_t out[8]) const {
        uint64_t cum[6]; read_cumulative(loc, cum_len, cum);
        for (int i = 0; i < 6; ++i) out[i] = cum[i];
        out[6] = out[7] = 0;                                  // codes 6,7 never occur in valid files (OOB write in the reference)
        uint64_t cur = 0;
        while (loc < end_pos) {
            uint8_t header = blocks_encoded_stream[loc++];
            int code = (header >> 5) & 0x7; uint64_t prefix = header & 0x1F, run_length;
            if (prefix < 31) run_length = prefix + 1;
            else run_length = 32 + bytecode_read(blocks_encoded_stream.data(), blocks_encoded_stream.size(), loc);
            if (cur + run_length > rel) { out[code] += (rel - cur); return; }
            out[code] += run_length; cur += run_length;
        }
    }

    // r-index.cpp:619-641
    rankvec rank_at_cached_encoded(uint64_t pos) const {
        if (!is_encoded()) return rank_at_cached(pos);
        rankvec r{}; uint64_t bid, bstart; locate_block(pos, bid, bstart);
        uint64_t nb = blocks_encoded_start_bits.size();
        if (bid >= nb) return r;
        size_t loc = (size_t) blocks_encoded_start_bits.get(bid);
        size_t end_pos = (bid + 1 < nb) ? (size_t) blocks_encoded_start_bits.get(bid + 1) : blocks_encoded_stream.size();
        uint64_t counts[8];
        ranks_at(loc, end_pos, pos - bstart, encoded_has_N ? 6 : 5, counts);
        for (size_t i = 0; i < C.size(); ++i) r[i] = counts[sym_map[NUC[i]]];
        return r;
    }

    // r-index.cpp:713-756 (encoded) and :1395-1428 (legacy) share this body; the rank
    // vectors differ. `a` is the reference's size_t symbol; bytes >= 0x80 would index
    // sym_map out of bounds there (sign-extended char) and are treated as absent here.
    bi_interval backward_extend_encoded(const bi_interval &bint, uint64_t a, uint64_t *distinct_blocks = nullptr) const {
        a &= 0xFF;
        uint64_t k = bint.forward, k_prime = bint.reverse; int64_t s = bint.size; size_t b = 0;
        rankvec rank_cache_ks = rank_at_cached_encoded(k + (uint64_t) s);
        rankvec rank_cache_k = rank_at_cached_encoded(k);
        if (distinct_blocks) {
            uint64_t b1, b2, t; locate_block(k + (uint64_t) s, b1, t); locate_block(k, b2, t);
            *distinct_blocks = (b1 == b2) ? 1 : 2;
        }
        const uint64_t comp_sym = complement(a);
        const uint64_t comp_idx = sym_map[comp_sym];
        while (b < 6 && sym_map[NUC[b]] < comp_idx) {
            const uint64_t idx = sym_map[complement(NUC[b])];
            k_prime += rank_cache_ks[idx] - rank_cache_k[idx];
            b++;
        }
        uint64_t rank_ks = rank_cache_ks[sym_map[a]], rank_k = rank_cache_k[sym_map[a]];
        if (rank_k >= rank_ks) return bi_interval{0, 0, 0};
        s = (int64_t) (rank_ks - rank_k);
        k = rank_k + C[sym_map[a]];
        return bi_interval{k, k_prime, s};
    }

    // ---- whole-read backward search: the query_tags path ----------------------
    // symbol_to_code, r-index.hpp:664-667 (bytes outside nuc alias code 0, the endmarker)
    static int symbol_to_code(uint64_t symbol) { for (int i = 0; i < 6; ++i) if ((uint64_t) NUC[i] == symbol) return i; return 0; }

    // FastLocate::rankAt (4-argument form), r-index.cpp:558-568: true rank of byte `symbol` at pos
    uint64_t rankAt(uint64_t pos, uint64_t symbol) const {
        uint64_t bid, bstart; locate_block(pos, bid, bstart);
        if (bid >= blocks.size()) return 0;
        const RunBlock &b = blocks[bid];
        uint64_t col = sym_map[symbol & 0xFF];
        return block_rankAt(b, pos - bstart, symbol) + (col < b.cum.size() ? b.cum[col] : 0);
    }

    // EncodedBlock::rank_of_code, r-index.cpp:114-131
    uint64_t rank_of_code(size_t &loc, size_t end_pos, int target_code, uint64_t rel) const {
        uint64_t rank = 0, cur = 0;
        while (loc < end_pos) {
            uint8_t header = blocks_encoded_stream[loc++];
            int code = (header >> 5) & 0x7; uint64_t prefix = header & 0x1F;
            uint64_t run_length = prefix < 31 ? prefix + 1 : 32 + bytecode_read(blocks_encoded_stream.data(), blocks_encoded_stream.size(), loc);
            if (code == target_code) {
                if (cur + run_length > rel) { rank += rel - cur; break; }
                rank += run_length;
            }
            cur += run_length;
            if (cur > rel) break;
        }
        return rank;
    }

    // FastLocate::rankAt_encoded, r-index.cpp:570-590. The reference constructs the block reader with SIX
    // cumulatives whatever the file holds (:577): on an encoded file without N (five cumulatives) the sixth read
    // swallows run-header bytes and the scan starts late. Restated literally; the product refuses that case.
    uint64_t rankAt_encoded(uint64_t pos, uint64_t symbol) const {
        if (!is_encoded()) return rankAt(pos, symbol);
        uint64_t bid, bstart; locate_block(pos, bid, bstart);
        uint64_t nb = blocks_encoded_start_bits.size();
        if (bid >= nb) return 0;
        size_t loc = (size_t) blocks_encoded_start_bits.get(bid);
        size_t end_pos = (bid + 1 < nb) ? (size_t) blocks_encoded_start_bits.get(bid + 1) : blocks_encoded_stream.size();
        uint64_t cum[6]; read_cumulative(loc, 6, cum);
        int target_code = symbol_to_code(symbol);
        return rank_of_code(loc, end_pos, target_code, pos - bstart) + cum[target_code];
    }

    // FastLocate::LF (range form), r-index.cpp:650-686, and LF_encoded, :689-711. A pattern byte reaches these as
    // size_t(char); bytes >= 0x80 would index sym_map out of bounds there and are taken modulo 256 here.
    std::pair<uint64_t, uint64_t> LF_encoded(std::pair<uint64_t, uint64_t> range, uint64_t sym) const {
        sym &= 0xFF;
        if (!is_encoded() && !sym_map[sym]) return {1, 0};                       // legacy LF only (:653)
        if (range.first > range.second) return {1, 0};
        range.first = rankAt_encoded(range.first, sym);
        uint64_t sym_inside = rankAt_encoded(range.second + 1, sym) - range.first;
        if (sym_inside == 0) return {1, 0};
        range.first += C[sym_map[sym]];
        range.second = range.first + sym_inside - 1;
        return range;
    }

    // FastLocate::count_encoded, r-index.hpp:550-556 (count :541-547 is the same loop over the legacy LF)
    std::pair<uint64_t, uint64_t> count_encoded(const char *pattern, size_t len) const {
        std::pair<uint64_t, uint64_t> range{0, bwt_size() - 1};
        for (size_t i = len; i > 0; --i) range = LF_encoded(range, (uint8_t) pattern[i - 1]);
        return range;
    }

    // ---- locate: toehold sample + phi ------------------------------------------------------
    static const uint64_t NO_POSITION = UINT64_MAX;                                  // r-index.hpp:49
    uint64_t getSample(uint64_t run_id) const { return samples.get(run_id); }        // r-index.hpp:616-618
    uint64_t seqId(uint64_t offset) const { return offset / max_length; }            // r-index.hpp:429
    uint64_t seqOffset(uint64_t offset) const { return offset % max_length; }        // r-index.hpp:431
    uint64_t locateFirst() const { return getSample(0); }                            // r-index.hpp:490-492

    // FastLocate::locateNext, r-index.cpp:1369-1372 (phi): the SA value of the row below the one holding prev.
    // Defined UB: no predecessor in `last` / a run id past the samples read as sample 0.
    uint64_t locateNext(uint64_t prev) const {
        uint64_t j = last.predecessor_index(prev);
        if (j == UINT64_MAX) return 0;
        uint64_t run = last_to_run.get(j) + 1;
        return (run < samples.size() ? samples.get(run) : 0) + (prev - last.ones[j]);
    }

    // The run lookup at the head of locate_encoded, r-index.cpp:1308-1329: logical run (block_id * block size + index in
    // the block) containing BWT row pos, and the row it starts at. skip_header reads SIX ByteCode values whatever the
    // file holds (:85-90): on an encoded file without N the scan starts inside the run headers (restated literally).
    void run_id_and_start(uint64_t pos, uint64_t &run_id, uint64_t &run_start) const {
        uint64_t bid, bstart; locate_block(pos, bid, bstart);
        uint64_t nb = blocks_encoded_start_bits.size();
        run_id = 0; run_start = 0;
        if (!is_encoded()) {
            // EXTENSION, not reference behaviour: the legacy twin FastLocate::locate has an empty body (:1281-1296) and
            // locate_encoded is undefined on a legacy file. The same lookup over the legacy block's runs, so that the
            // shipped legacy files (the only ones with samples the reference ships) can be located on.
            if (bid >= blocks.size()) return;
            uint64_t cur = 0, runnum = 0;
            for (const auto &rn : blocks[bid].runs) { if (cur + rn.second > pos - bstart) break; cur += rn.second; runnum++; }
            run_id = bid * 10 + runnum; run_start = bstart + cur;
            return;
        }
        if (bid >= nb) return;
        size_t loc = (size_t) blocks_encoded_start_bits.get(bid);
        for (int i = 0; i < 6; ++i) (void) bytecode_read(blocks_encoded_stream.data(), blocks_encoded_stream.size(), loc);
        uint64_t cur = 0, runnum = 0;
        for (;;) {
            if (loc >= blocks_encoded_stream.size()) break;                              // reference: unbounded loop; defined as "last run seen"
            uint8_t header = blocks_encoded_stream[loc++];
            uint64_t prefix = header & 0x1F;
            uint64_t run_length = prefix < 31 ? prefix + 1 : 32 + bytecode_read(blocks_encoded_stream.data(), blocks_encoded_stream.size(), loc);
            if (cur + run_length > pos - bstart) break;
            cur += run_length; runnum++;
        }
        run_id = bid * (encoded_block_size ? encoded_block_size : 10) + runnum;
        run_start = bstart + cur;
    }

    // FastLocate::locate_encoded, r-index.cpp:1300-1343: sorted unique SEQUENCE IDS of the rows of [state.first,
    // state.second]. `sa` (optional) receives the raw SA values row by row (the values the reference holds in `first`).
    std::vector<uint64_t> locate_encoded(std::pair<uint64_t, uint64_t> state, uint64_t first = NO_POSITION, std::vector<uint64_t> *sa = nullptr) const {
        std::vector<uint64_t> result;
        if (state.second < state.first) return result;
        uint64_t offset_of_first = state.first;
        if (first == NO_POSITION) {
            uint64_t run_id; run_id_and_start(state.first, run_id, offset_of_first);
            first = getSample(run_id);
        }
        while (offset_of_first < state.first) { first = locateNext(first); offset_of_first++; }
        result.push_back(seqId(first)); if (sa) sa->push_back(first);
        for (uint64_t i = state.first + 1; i <= state.second; i++) {
            first = locateNext(first);
            result.push_back(seqId(first)); if (sa) sa->push_back(first);
        }
        std::sort(result.begin(), result.end());
        result.erase(std::unique(result.begin(), result.end()), result.end());
        return result;
    }

    // r-index.cpp:758-764 / :1500-1509
    bi_interval forward_extend_encoded(const bi_interval &bint, uint64_t symbol, uint64_t *distinct_blocks = nullptr) const {
        bi_interval tmp{bint.reverse, bint.forward, bint.size};
        uint64_t comp = complement(symbol & 0xFF);
        tmp = backward_extend_encoded(tmp, comp, distinct_blocks);
        return bi_interval{tmp.reverse, tmp.forward, tmp.size};
    }
};

// ---------------------------------------------------------------------------------------------
// TagArray (query side). tag_arrays.hpp:116-120: length_bits = 9, encoded_start_every_k_run = 10.
struct TagArray {
    enum Layout { SDSL_COMPACT = 0, BYTECODE_LEN = 1 };
    int layout = SDSL_COMPACT;
    static const unsigned length_bits = 9;
    static const uint64_t encoded_start_every_k_run = 10;
    IntVec encoded_runs_iv;                 // sdsl-compact items
    std::vector<uint8_t> encoded_runs;      // ByteCode stream
    SdVec encoded_runs_starts_sd, bwt_intervals;

    // tag_arrays.cpp:766-776
    void load_compressed_tags_compact(Cursor &c) {
        layout = SDSL_COMPACT;
        encoded_runs_iv.load(c, 0); encoded_runs_starts_sd.load(c); bwt_intervals.load(c);
    }
    // tag_arrays.cpp:735-763
    void load_compressed_tags(Cursor &c) {
        layout = BYTECODE_LEN;
        uint64_t nbytes = c.u64();
        const uint8_t *src = c.bytes(nbytes);
        encoded_runs.assign(src, src + nbytes);
        encoded_runs_starts_sd.load(c); bwt_intervals.load(c);
    }

    // gbwtgraph::Position::encode(make_pos_t(id, rev, off)) = id<<11 | rev<<10 | off&0x3FF
    // (mirrors tag_arrays.cpp:38-45; the mapping itself is un-vendored -> "parity unpinned" for it)
    static uint64_t position_encode(uint64_t node, uint64_t rev, uint64_t off) { return (node << 11) | ((rev & 1) << 10) | (off & 0x3FF); }

    // Shared body of query_compressed (tag_arrays.cpp:780-854) and query_compressed_compact (:856-890).
    // Fills `out` with the sorted unique values; returns number_of_runs. `skipped` = items advanced over.
    // Defined UB: select beyond the last sampled start -> end of the item array; items / bytes past the
    // end read as 0.
    uint64_t query(uint64_t start, uint64_t end, std::vector<uint64_t> &out, uint64_t *skipped = nullptr) const {
        uint64_t first_bit_index = bwt_intervals.rank1(start + 1);
        uint64_t end_bit_index = bwt_intervals.rank1(end + 1);
        uint64_t run_nums = end_bit_index - first_bit_index + 1;
        uint64_t number_of_runs = run_nums;
        out.clear();
        uint64_t current_tag_run_index = first_bit_index - (first_bit_index % encoded_start_every_k_run);
        uint64_t move_tags = first_bit_index % encoded_start_every_k_run;
        uint64_t sel = encoded_runs_starts_sd.select1(current_tag_run_index / encoded_start_every_k_run + 1);
        if (skipped) *skipped = move_tags > 1 ? move_tags - 1 : 0;
        if ((int64_t) run_nums <= 0) return number_of_runs;   // reference: size_t underflow -> unbounded loop; defined as empty
        out.reserve(run_nums);
        if (layout == SDSL_COMPACT) {
            uint64_t n_items = encoded_runs_iv.size();
            uint64_t item_index = (sel == UINT64_MAX) ? n_items : sel;
            while (move_tags > 1) { item_index++; move_tags--; }
            while (run_nums > 0) {
                uint64_t enc = (item_index < n_items) ? encoded_runs_iv.get(item_index) : 0; item_index++;
                out.push_back(position_encode(enc >> 11, (enc >> 10) & 1, enc & 0x3FF));      // decode_run_length_compact :47-55
                run_nums--;
            }
        } else {
            size_t loc = (sel == UINT64_MAX) ? encoded_runs.size() : (size_t) sel;
            while (move_tags > 1) { (void) bytecode_read(encoded_runs.data(), encoded_runs.size(), loc); move_tags--; }
            while (run_nums > 0) {
                uint64_t enc = bytecode_read(encoded_runs.data(), encoded_runs.size(), loc);
                out.push_back(position_encode(enc >> (11 + length_bits), (enc >> 10) & 1, enc & 0x3FF)); // decode_run :59-70
                run_nums--;
            }
        }
        std::sort(out.begin(), out.end());
        out.erase(std::unique(out.begin(), out.end()), out.end());
        return number_of_runs;
    }
};

} // namespace orc
