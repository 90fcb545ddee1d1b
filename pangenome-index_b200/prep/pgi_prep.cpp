// Data-prep tools (CPU): everything needed to get from sequences to the files find_mems loads,
// without SDSL / GBWT / grlBWT:
//   * multi-string BWT (S1 $1 S2 $2 ... with $1<$2<...<'A', every $ written '\n';
//     convention pinned by /root/reference tests/test_rindex.cpp:35-99) -> grlBWT .rl_bwt records
//   * r-index construction = restatement of FastLocate::FastLocate(std::string)
//     (src/r-index.cpp:778-1139, calculate_C include/pangenome_index/r-index.hpp:440-482)
//   * FastLocate::serialize (legacy, :266-293) and serialize_encoded (:297-376)
//   * tag arrays in the sdsl-compact layout written by merge_tags (src/tag_arrays.cpp:622-654,940-974)
//     and in the ByteCode(+9 length bits) layout read by load_compressed_tags (:735-763)
//   * read sampling with substitution / indel errors.
// These are tools around the hot path, not the hot path.
#include "sais.hpp"
#include "sdsl_writer.hpp"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>

using namespace pgi_prep;

namespace {

thread_local std::string g_err;
const unsigned char NUC[6] = {'\n', 'A', 'C', 'G', 'N', 'T'};

struct Run { uint8_t sym; uint64_t len; };

uint8_t *dup_out(const std::vector<uint8_t> &v, uint64_t *len) {
    uint8_t *p = (uint8_t *) std::malloc(v.size() + 1);
    if (!v.empty()) std::memcpy(p, v.data(), v.size());
    *len = v.size();
    return p;
}

// ---- .rl_bwt (deps/grlBWT/include/bwt_io.h:40-58,96-135): u64 sb, u64 fb, then (sym: sb bytes, len: fb bytes)
std::vector<Run> parse_rlbwt(const uint8_t *d, uint64_t len) {
    if (len < 16) throw std::runtime_error("rl_bwt: too short");
    uint64_t sb, fb; std::memcpy(&sb, d, 8); std::memcpy(&fb, d + 8, 8);
    if (sb == 0 || sb > 8 || fb == 0 || fb > 8) throw std::runtime_error("rl_bwt: bad header");
    uint64_t bpr = sb + fb, nrec = (len - 16) / bpr;
    std::vector<Run> runs; runs.reserve(nrec);
    for (uint64_t i = 0; i < nrec; ++i) {
        uint64_t sym = 0, freq = 0;
        std::memcpy(&sym, d + 16 + i * bpr, sb); std::memcpy(&freq, d + 16 + i * bpr + sb, fb);
        runs.push_back(Run{(uint8_t) sym, freq});
    }
    return runs;
}

std::vector<uint8_t> write_rlbwt(const std::vector<Run> &runs) {
    uint64_t maxlen = 1; for (auto &r : runs) maxlen = std::max(maxlen, r.len);
    uint64_t fb = (bits_hi(maxlen) + 8) / 8, sb = 1;
    ByteSink out; out.u64(sb); out.u64(fb);
    for (auto &r : runs) { out.u8(r.sym); out.bytes(&r.len, fb); }
    return out.buf;
}

// ---- BWT of newline-terminated sequences via SA-IS
// text: sequences each terminated by '\n'. Returns BWT bytes (n = len) and optionally the suffix array.
void build_bwt(const uint8_t *text, uint64_t n, std::vector<uint8_t> &bwt, std::vector<int64_t> *sa_out) {
    if (n == 0 || text[n - 1] != '\n') throw std::runtime_error("build_bwt: text must end with a newline");
    uint64_t nseq = 0; for (uint64_t i = 0; i < n; ++i) nseq += text[i] == '\n';
    bwt.resize(n);
    auto emit = [&](auto *SA) {
        for (uint64_t i = 0; i < n; ++i) {
            uint64_t p = (uint64_t) SA[i + 1];
            bwt[i] = p == 0 ? text[n - 1] : text[p - 1];
            if (sa_out) (*sa_out)[i] = (int64_t) p;
        }
    };
    if (sa_out) sa_out->resize(n);
    if (n + 1 < 0x7FFFFFF0ULL && nseq + 300 < 0x7FFFFFF0ULL) {
        std::vector<int32_t> s(n + 1), SA(n + 1);
        int32_t k = 0;
        for (uint64_t i = 0; i < n; ++i) s[i] = text[i] == '\n' ? ++k : (int32_t) (nseq + 1 + text[i]);
        s[n] = 0;
        SAIS<int32_t>::run(s.data(), SA.data(), (int32_t) (n + 1), (int32_t) (nseq + 1 + 256));
        emit(SA.data());
    } else {
        std::vector<int64_t> s(n + 1), SA(n + 1);
        int64_t k = 0;
        for (uint64_t i = 0; i < n; ++i) s[i] = text[i] == '\n' ? ++k : (int64_t) (nseq + 1 + text[i]);
        s[n] = 0;
        SAIS<int64_t>::run(s.data(), SA.data(), (int64_t) (n + 1), (int64_t) (nseq + 1 + 256));
        emit(SA.data());
    }
}

std::vector<Run> runs_of(const std::vector<uint8_t> &bwt) {
    std::vector<Run> runs;
    for (uint64_t i = 0; i < bwt.size();) {
        uint64_t j = i; while (j < bwt.size() && bwt[j] == bwt[i]) ++j;
        runs.push_back(Run{bwt[i], j - i}); i = j;
    }
    return runs;
}

// ---- r-index construction (restating FastLocate::FastLocate(std::string))
struct RIndex {
    uint64_t n = 0, n_seq = 0, max_length = 1, total_runs = 0;
    uint8_t sym_map[256]; std::vector<uint64_t> C;
    std::vector<Run> lruns;                       // logical runs (every endmarker its own run)
    std::vector<uint64_t> lrun_start;             // first BWT row of each logical run
    uint64_t n_blocks_serialized = 0;
    std::vector<uint64_t> block_starts;           // non-empty blocks only
    std::vector<std::vector<uint64_t>> block_cum; // per serialized block
    std::vector<uint64_t> samples, last_ones, last_to_run;
    std::vector<uint64_t> row_textpos;            // optional: packed (seq_id * max_length + seq_offset) per BWT row
    static const uint64_t block_size = 10;
};

void build_rindex(const std::vector<Run> &records, RIndex &R, bool keep_row_textpos, bool with_samples) {
    // calculate_C (r-index.hpp:440-482)
    uint64_t freqs[256] = {0}; R.n = 0;
    for (auto &r : records) { freqs[r.sym] += r.len; R.n += r.len; }
    std::memset(R.sym_map, 0, 256);
    R.C.clear(); uint64_t k = 0, acc = 0;
    for (int i = 0; i < 256; ++i) if (freqs[i]) { R.sym_map[i] = (uint8_t) k++; R.C.push_back(acc); acc += freqs[i]; }
    const uint64_t sigma = R.C.size();
    R.n_seq = sigma > 1 ? R.C[1] - R.C[0] : R.n;                    // tot_strings(), r-index.hpp:484-486
    // logical runs + blocks (r-index.cpp:796-958)
    R.lruns.clear();
    for (auto &r : records) {
        if (r.sym == '\n') for (uint64_t i = 0; i < r.len; ++i) R.lruns.push_back(Run{r.sym, 1});
        else R.lruns.push_back(r);
    }
    R.total_runs = R.lruns.size();
    R.n_blocks_serialized = R.total_runs / R.block_size + 1;
    R.block_starts.clear(); R.block_cum.assign(R.n_blocks_serialized, std::vector<uint64_t>(8, 0));
    R.lrun_start.resize(R.total_runs);
    std::vector<uint64_t> cum(sigma, 0); uint64_t pos = 0;
    for (uint64_t i = 0; i < R.total_runs; ++i) {
        if (i % R.block_size == 0) { R.block_cum[i / R.block_size] = cum; R.block_starts.push_back(pos); }
        R.lrun_start[i] = pos;
        cum[R.sym_map[R.lruns[i].sym]] += R.lruns[i].len; pos += R.lruns[i].len;
    }
    R.samples.clear(); R.last_ones.clear(); R.last_to_run.clear(); R.row_textpos.clear();
    if (!with_samples && !keep_row_textpos) return;
    // plain BWT + LF + run id per row (the reference walks psi()/rankAt per step, r-index.cpp:1025-1094)
    const uint64_t n = R.n;
    std::vector<uint8_t> bwt(n); std::vector<uint32_t> row_run;
    bool small_runs = R.total_runs < 0xFFFFFFFFULL;
    if (!small_runs) throw std::runtime_error("build_rindex: too many runs for the sample walk");
    row_run.resize(n);
    for (uint64_t i = 0; i < R.total_runs; ++i)
        for (uint64_t j = 0; j < R.lruns[i].len; ++j) { bwt[R.lrun_start[i] + j] = R.lruns[i].sym; row_run[R.lrun_start[i] + j] = (uint32_t) i; }
    std::vector<uint64_t> LF(n); { std::vector<uint64_t> cnt(R.C); for (uint64_t i = 0; i < n; ++i) LF[i] = cnt[R.sym_map[bwt[i]]]++; }
    struct Rec { uint64_t seq_id, seq_offset, run_id; };
    std::vector<Rec> heads, tails;
    std::vector<std::pair<uint64_t, uint64_t>> row_pos;   // (row, t) per sequence, resolved after the walk
    std::vector<uint64_t> seq_len(R.n_seq, 0);
    std::vector<uint32_t> row_seq; std::vector<uint64_t> row_t;
    if (keep_row_textpos) { row_seq.assign(n, 0); row_t.assign(n, 0); }
    for (uint64_t i = 0; i < R.n_seq; ++i) {
        size_t h0 = heads.size(), t0 = tails.size();
        uint64_t seq_offset = 0, run_id = row_run[i];
        if (i == 0 || run_id != row_run[i - 1]) heads.push_back(Rec{i, seq_offset, run_id});
        if (i + 1 >= R.n_seq || run_id != row_run[i + 1]) tails.push_back(Rec{i, seq_offset, run_id});
        if (keep_row_textpos) { row_seq[i] = (uint32_t) i; row_t[i] = 0; }
        uint8_t cur_sym = bwt[i]; uint64_t cur_row = LF[i];
        seq_offset++;
        while (cur_sym != '\n') {
            uint64_t rid = row_run[cur_row];
            uint64_t rfirst = R.lrun_start[rid], rlast = rfirst + R.lruns[rid].len - 1;
            if (cur_row == rfirst) heads.push_back(Rec{i, seq_offset, rid});
            if (cur_row == rlast) tails.push_back(Rec{i, seq_offset, rid});
            if (keep_row_textpos) { row_seq[cur_row] = (uint32_t) i; row_t[cur_row] = seq_offset; }
            cur_sym = bwt[cur_row]; cur_row = LF[cur_row];
            seq_offset++;
        }
        for (size_t j = h0; j < heads.size(); ++j) heads[j].seq_offset = seq_offset - 1 - heads[j].seq_offset;
        for (size_t j = t0; j < tails.size(); ++j) tails[j].seq_offset = seq_offset - 1 - tails[j].seq_offset;
        seq_len[i] = seq_offset;
        R.max_length = std::max(R.max_length, seq_offset);
    }
    if (keep_row_textpos) {
        R.row_textpos.resize(n);
        for (uint64_t r = 0; r < n; ++r) R.row_textpos[r] = (uint64_t) row_seq[r] * R.max_length + (seq_len[row_seq[r]] - 1 - row_t[r]);
    }
    if (!with_samples) return;
    std::stable_sort(heads.begin(), heads.end(), [](const Rec &a, const Rec &b) { return a.run_id < b.run_id; });
    std::stable_sort(tails.begin(), tails.end(), [](const Rec &a, const Rec &b) {
        return a.seq_id < b.seq_id || (a.seq_id == b.seq_id && a.seq_offset < b.seq_offset); });
    if (heads.size() < R.total_runs || tails.size() < R.total_runs) throw std::runtime_error("build_rindex: missing samples (BWT is not a valid multi-string BWT)");
    R.samples.resize(R.total_runs); R.last_ones.resize(R.total_runs); R.last_to_run.resize(R.total_runs);
    for (uint64_t i = 0; i < R.total_runs; ++i) {
        R.samples[i] = heads[i].seq_id * R.max_length + heads[i].seq_offset;
        R.last_ones[i] = tails[i].seq_id * R.max_length + tails[i].seq_offset;
        R.last_to_run[i] = tails[i].run_id;
    }
}

void serialize_common(ByteSink &out, const RIndex &R, uint64_t flags, bool with_samples) {
    out.u32(0x6B3741D8u); out.u32(1); out.u64(R.max_length); out.u64(flags);
    if (with_samples) {
        PackedVec s(R.total_runs, bits_length((R.n_seq - 1) * R.max_length + (R.max_length - 1)));
        for (uint64_t i = 0; i < R.total_runs; ++i) s.set(i, R.samples[i]);
        s.serialize_iv0(out);
        write_sd_vector(out, R.n_seq * R.max_length, R.last_ones);
        unsigned w = bits_length(R.total_runs - 1); if (w == 0) w = 64;
        PackedVec l(R.total_runs, w);
        for (uint64_t i = 0; i < R.total_runs; ++i) l.set(i, R.last_to_run[i]);
        l.serialize_iv0(out);
    } else {
        // find_mems never reads samples / last / last_to_run: written empty.
        PackedVec s(0, 64); s.serialize_iv0(out);
        write_sd_vector(out, 0, std::vector<uint64_t>());
        s.serialize_iv0(out);
    }
    PackedVec sm(256, 8); for (int i = 0; i < 256; ++i) sm.set(i, R.sym_map[i]); sm.serialize_fixed(out);
    PackedVec Cv(R.C.size(), 64); for (size_t i = 0; i < R.C.size(); ++i) Cv.set(i, R.C[i]); Cv.serialize_fixed(out);
    write_sd_vector(out, R.n, R.block_starts);
    out.u64(R.n);
}

std::vector<uint8_t> serialize_legacy(const RIndex &R, bool with_samples) {
    ByteSink out; serialize_common(out, R, 0, with_samples);
    out.u64(R.n_blocks_serialized);
    for (uint64_t b = 0; b < R.n_blocks_serialized; ++b) {
        const auto &cum = R.block_cum[b];
        PackedVec cv(cum.size(), 64); for (size_t i = 0; i < cum.size(); ++i) cv.set(i, cum[i]); cv.serialize_fixed(out);
        uint64_t r0 = b * R.block_size, r1 = std::min(R.total_runs, r0 + R.block_size);
        uint64_t nr = r1 > r0 ? r1 - r0 : 0; out.u64(nr);
        for (uint64_t r = r0; r < r1; ++r) { out.u64(R.lruns[r].sym); out.u64(R.lruns[r].len); }
    }
    return out.buf;
}

std::vector<uint8_t> serialize_encoded(const RIndex &R, bool with_samples) {
    ByteSink out; serialize_common(out, R, 1, with_samples);
    out.u64(R.block_size);
    bool hasN = false; for (auto &r : R.lruns) if (r.sym == 'N') { hasN = true; break; }
    out.u8(hasN ? 1 : 0);
    std::vector<uint8_t> stream; std::vector<uint64_t> offs;
    for (uint64_t b = 0; b < R.n_blocks_serialized; ++b) {
        offs.push_back(stream.size());
        for (uint64_t v : R.block_cum[b]) bytecode_write(stream, v);
        uint64_t r0 = b * R.block_size, r1 = std::min(R.total_runs, r0 + R.block_size);
        for (uint64_t r = r0; r < r1; ++r) {
            int code = 0; for (int i = 0; i < 6; ++i) if (NUC[i] == R.lruns[r].sym) { code = i; break; }
            uint64_t prefix = std::min<uint64_t>(R.lruns[r].len - 1, 31);
            stream.push_back((uint8_t) (((code & 7) << 5) | (prefix & 0x1F)));
            if (prefix == 31) bytecode_write(stream, R.lruns[r].len - 32);
        }
    }
    unsigned w = bits_length(offs.empty() ? 0 : offs.back()); if (w == 0) w = 64;
    PackedVec ov(offs.size(), w); for (size_t i = 0; i < offs.size(); ++i) ov.set(i, offs[i]); ov.serialize_iv0(out);
    out.u64(stream.size()); out.bytes(stream.data(), stream.size());
    return out.buf;
}

// ---- tags. row_value[r] = encoded graph position (node<<11 | rev<<10 | off) of BWT row r; rows of
// endmarker suffixes carry 0 (merge_tags.cpp:619-622). Runs longer than 511 are split (tag_arrays.cpp:942-957).
struct TagRuns { std::vector<uint64_t> value, len; };
TagRuns tag_runs_of(const uint64_t *row_value, uint64_t n) {
    TagRuns t;
    for (uint64_t i = 0; i < n;) {
        uint64_t j = i; while (j < n && row_value[j] == row_value[i]) ++j;
        uint64_t l = j - i;
        while (l >= 512) { t.value.push_back(row_value[i]); t.len.push_back(511); l -= 511; }
        if (l > 0) { t.value.push_back(row_value[i]); t.len.push_back(l); }
        i = j;
    }
    return t;
}
// layout 0: sdsl-compact [int_vector<0> items][sd starts (item index of runs 0,10,..)][sd bwt_intervals]
// layout 1: ByteCode [u64 nbytes][ByteCode off|rev<<10|len<<11|node<<20][sd starts (byte offsets)][sd bwt_intervals]
std::vector<uint8_t> serialize_tags(const TagRuns &t, int layout) {
    ByteSink out; uint64_t nr = t.value.size();
    std::vector<uint64_t> starts, bwt_ones; uint64_t cumpos = 0;
    if (layout == 0) {
        uint64_t max_node = 0; for (uint64_t v : t.value) max_node = std::max(max_node, v >> 11);
        unsigned width = 11 + bits_hi(max_node) + 1;
        PackedVec items(nr, width);
        for (uint64_t i = 0; i < nr; ++i) {
            items.set(i, t.value[i]);
            if (i % 10 == 0) starts.push_back(i);
            bwt_ones.push_back(cumpos); cumpos += t.len[i];
        }
        items.serialize_iv0(out);
    } else {
        std::vector<uint8_t> stream;
        for (uint64_t i = 0; i < nr; ++i) {
            if (i % 10 == 0) starts.push_back(stream.size());
            uint64_t v = t.value[i];
            uint64_t enc = (v & 0x3FF) | (((v >> 10) & 1) << 10) | ((t.len[i] & 0x1FF) << 11) | ((v >> 11) << 20);
            bytecode_write(stream, enc);
            bwt_ones.push_back(cumpos); cumpos += t.len[i];
        }
        out.u64(stream.size()); out.bytes(stream.data(), stream.size());
    }
    write_sd_vector(out, (starts.empty() ? 0 : starts.back()) + 1, starts);
    write_sd_vector(out, cumpos + 1, bwt_ones);
    return out.buf;
}

inline uint64_t splitmix(uint64_t &s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL; z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL; return z ^ (z >> 31);
}

} // namespace

extern "C" {

const char *prep_last_error() { return g_err.c_str(); }
void prep_free(void *p) { std::free(p); }

// text (newline-terminated sequences) -> .rl_bwt file bytes; optionally also the plain BWT.
int prep_bwt_from_text(const uint8_t *text, uint64_t n, uint8_t **rlbwt, uint64_t *rlbwt_len, uint8_t **bwt_plain) {
    try {
        std::vector<uint8_t> bwt; build_bwt(text, n, bwt, nullptr);
        std::vector<uint8_t> f = write_rlbwt(runs_of(bwt));
        *rlbwt = dup_out(f, rlbwt_len);
        if (bwt_plain) { uint64_t l; *bwt_plain = dup_out(bwt, &l); }
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

// .rl_bwt bytes -> .ri bytes. encoded: 0 legacy layout, 1 ENCODED_BLOCKS. with_samples: 0 writes empty
// samples/last/last_to_run (enough for find_mems). row_textpos (optional, n entries, caller-allocated):
// packed seq_id*max_length+seq_offset of every BWT row; info[0..7] = n, n_seq, max_length, total_runs, sigma.
int prep_rindex_from_rlbwt(const uint8_t *rlbwt, uint64_t len, int encoded, int with_samples,
                           uint8_t **ri, uint64_t *ri_len, uint64_t *row_textpos, uint64_t *info) {
    try {
        RIndex R; build_rindex(parse_rlbwt(rlbwt, len), R, row_textpos != nullptr, with_samples != 0);
        std::vector<uint8_t> f = encoded ? serialize_encoded(R, with_samples != 0) : serialize_legacy(R, with_samples != 0);
        *ri = dup_out(f, ri_len);
        if (row_textpos) std::memcpy(row_textpos, R.row_textpos.data(), R.n * 8);
        if (info) { info[0] = R.n; info[1] = R.n_seq; info[2] = R.max_length; info[3] = R.total_runs; info[4] = R.C.size(); }
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

// per-row encoded graph positions -> tags file bytes (layout 0 sdsl-compact, 1 ByteCode+len9)
int prep_tags_from_rows(const uint64_t *row_value, uint64_t n, int layout, uint8_t **out, uint64_t *out_len, uint64_t *n_runs) {
    try {
        TagRuns t = tag_runs_of(row_value, n);
        std::vector<uint8_t> f = serialize_tags(t, layout);
        *out = dup_out(f, out_len); if (n_runs) *n_runs = t.value.size();
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

// (value, length) tag runs -> tags file bytes; runs longer than 511 are split (tag_arrays.cpp:942-957).
int prep_tags_from_runs(const uint64_t *value, const uint64_t *len, uint64_t n_in, int layout, uint8_t **out, uint64_t *out_len, uint64_t *n_runs) {
    try {
        TagRuns t; t.value.reserve(n_in + n_in / 16); t.len.reserve(n_in + n_in / 16);
        for (uint64_t i = 0; i < n_in; ++i) {
            uint64_t l = len[i];
            while (l >= 512) { t.value.push_back(value[i]); t.len.push_back(511); l -= 511; }
            if (l > 0) { t.value.push_back(value[i]); t.len.push_back(l); }
        }
        std::vector<uint8_t> f = serialize_tags(t, layout);
        *out = dup_out(f, out_len); if (n_runs) *n_runs = t.value.size();
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

// Sample n_reads reads of read_len from newline-terminated sequences, uniform over valid start positions,
// with i.i.d. substitution rate `sub` and 1-bp indel rate `indel` (each split evenly insert/delete).
// Output: fixed-stride buffer reads[n_reads*read_len] (reads never span a newline).
int prep_sample_reads(const uint8_t *text, uint64_t n, uint64_t n_reads, uint64_t read_len, double sub, double indel,
                      uint64_t seed, uint8_t *reads) {
    try {
        std::vector<uint64_t> st, ln; uint64_t s = 0, total = 0; const uint64_t margin = indel > 0 ? read_len / 8 + 16 : 0;
        for (uint64_t i = 0; i < n; ++i) if (text[i] == '\n') { if (i - s >= read_len + margin) { st.push_back(s); ln.push_back(i - s - read_len - margin + 1); total += ln.back(); } s = i + 1; }
        if (total == 0) throw std::runtime_error("sample_reads: no sequence long enough");
        std::vector<uint64_t> cum(st.size() + 1, 0); for (size_t i = 0; i < st.size(); ++i) cum[i + 1] = cum[i] + ln[i];
        const char ACGT[4] = {'A', 'C', 'G', 'T'};
        const uint64_t sub_t = (uint64_t) (sub * 18446744073709551615.0), ind_t = (uint64_t) (indel * 18446744073709551615.0);
#pragma omp parallel for schedule(static)
        for (int64_t r = 0; r < (int64_t) n_reads; ++r) {
            uint64_t rs = seed * 0x9E3779B97F4A7C15ULL + (uint64_t) r * 0xD1B54A32D192ED03ULL + 1;
            uint64_t u = splitmix(rs) % total;
            size_t k = (size_t) (std::upper_bound(cum.begin(), cum.end(), u) - cum.begin()) - 1;
            const uint8_t *src = text + st[k] + (u - cum[k]);
            uint8_t *dst = reads + (uint64_t) r * read_len;
            uint64_t si = 0;
            for (uint64_t j = 0; j < read_len; ++j) {
                if (ind_t && splitmix(rs) < ind_t) {
                    if (splitmix(rs) & 1) { dst[j] = (uint8_t) ACGT[splitmix(rs) & 3]; continue; }   // insertion
                    else si++;                                                                      // deletion
                }
                if (si >= read_len + margin) si = read_len + margin - 1;
                uint8_t c = src[si++];
                if (sub_t && splitmix(rs) < sub_t) { uint8_t d; do { d = (uint8_t) ACGT[splitmix(rs) & 3]; } while (d == c); c = d; }
                dst[j] = c;
            }
        }
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

} // extern "C"
