// panindexer.hpp — C++ host facade over the C ABI of include/pgi.h, with the reference's class and function
// names so that code written against parsaeskandar/pangenome-index's find_mems path compiles against this
// header and links libpgi.so instead of libpanindexer.a + SDSL + GBWT.
//
// Mirrors (reference file:line):
//   panindexer::FastLocate            include/pangenome_index/r-index.hpp:44-714   (query side only)
//     load / load_encoded             src/r-index.cpp:378-459
//     backward/forward_extend[_encoded]   src/r-index.cpp:713-764, 1395-1428, 1500-1509
//     count / count_encoded           include/pangenome_index/r-index.hpp:541-556 (LF / LF_encoded src/r-index.cpp:650-711)
//     locate / locate_encoded, decompressSA / decompressDA, seqId / seqOffset   src/r-index.cpp:1300-1372
//     bwt_size, is_encoded, complement, sym_map, C
//   panindexer::TagArray              include/pangenome_index/tag_arrays.hpp:27-128 (query side only)
//     load_compressed_tags[_compact]  src/tag_arrays.cpp:739-776
//     query_compressed[_compact]      src/tag_arrays.cpp:780-890   (prints to std::cout like the reference)
//   panindexer::MEM, find_all_mems    include/pangenome_index/algorithm.hpp:644-757
// Error behaviour: the reference throws sdsl::simple_sds::InvalidData on bad files; here std::runtime_error with
// the message of pgi_last_error(). Every call runs on the GPU; without a CUDA device construction throws.
//
// The per-read / per-extension methods exist for drop-in compatibility and tests; each is one small GPU launch.
// Throughput comes from the batched entry points (find_all_mems_batch / pgi_find_mems), which is what the
// find_mems driver in this directory uses.
#pragma once
#include <array>
#include <cstdint>
#include <iostream>
#include <istream>
#include <iterator>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>
#include "../../include/pgi.h"

namespace panindexer {

static const unsigned char NENDMARKER = '\n';                               // utils.hpp:9

inline void pgi_check(int rc) { if (rc != PGI_OK) throw std::runtime_error(std::string("pgi: ") + pgi_last_error()); }

inline std::string slurp(std::istream &in) {
    return std::string(std::istreambuf_iterator<char>(in), std::istreambuf_iterator<char>());
}

class FastLocate {
public:
    struct bi_interval {                                                     // r-index.hpp:118-130
        size_t forward, reverse; int64_t size;
        bi_interval() : forward(0), reverse(0), size(0) {}
        bi_interval(size_t f, size_t r, int64_t s) : forward(f), reverse(r), size(s) {}
    };
    std::array<uint8_t, 256> sym_map{};
    std::vector<size_t> C;

    FastLocate() = default;
    explicit FastLocate(int device) : device_(device) {}
    FastLocate(const FastLocate &) = delete;
    FastLocate &operator=(const FastLocate &) = delete;
    ~FastLocate() { if (ix_) pgi_index_free(ix_); }

    // Both layouts are recognised from the header flag, exactly like load_encoded's fallback (r-index.cpp:422-427).
    void load_encoded(std::istream &in) { bytes_ = slurp(in); create(); }
    void load(std::istream &in) { load_encoded(in); }

    size_t bwt_size() const { return info_.n; }
    bool is_encoded() const { return info_.is_encoded != 0; }
    size_t complement(size_t symbol) const {
        switch (symbol) { case 'A': return 'T'; case 'C': return 'G'; case 'G': return 'C'; case 'T': return 'A';
                          case 'a': return 't'; case 'c': return 'g'; case 'g': return 'c'; case 't': return 'a'; default: return symbol; }
    }
    bi_interval backward_extend_encoded(const bi_interval &bint, size_t symbol) { return extend(bint, symbol, 0); }
    bi_interval forward_extend_encoded(const bi_interval &bint, size_t symbol) { return extend(bint, symbol, 1); }
    bi_interval backward_extend(const bi_interval &bint, size_t symbol) { return extend(bint, symbol, 0); }
    bi_interval forward_extend(const bi_interval &bint, size_t symbol) { return extend(bint, symbol, 1); }

    // Whole-read backward search, r-index.hpp:541-556: {first, second} of gbwt::range_type; no match = {1, 0}.
    // (count and count_encoded resolve to the same call: the library follows the file's own layout, as
    // src/query_tags.cpp:90-96 does by hand.)
    typedef std::pair<size_t, size_t> range_type;
    range_type count_encoded(std::string &pattern) const {
        if (!ix_) throw std::runtime_error("FastLocate: no index loaded");
        const uint64_t off[2] = {0, (uint64_t) pattern.size()};
        pgi_result r;
        pgi_check(pgi_query_tags(ix_, pattern.data(), off, 1, nullptr, &r));
        range_type out((size_t) r.mems[0].bwt_start, (size_t) (r.mems[0].bwt_start + (uint64_t) r.mems[0].size - 1));
        pgi_result_free(&r);
        return out;
    }
    range_type count(std::string &pattern) const { return count_encoded(pattern); }

    // FastLocate::locate_encoded, src/r-index.cpp:1300-1343: sorted unique sequence ids of the rows of `state`
    // (toehold sample of the first row's run + one phi step per row). `first` must be NO_POSITION (the library finds the
    // toehold itself). locate() forwards to it (the reference's legacy body is empty, :1281-1296).
    static constexpr size_t NO_POSITION = ~(size_t) 0;                       // r-index.hpp:49
    std::vector<size_t> locate_encoded(range_type state, size_t first = NO_POSITION) const {
        if (!ix_) throw std::runtime_error("FastLocate: no index loaded");
        if (first != NO_POSITION) throw std::runtime_error("FastLocate::locate_encoded: a caller-provided toehold is not supported");
        const uint64_t a = state.first, b = state.second; pgi_locate_result r;
        pgi_check(pgi_locate_batch(ix_, 1, &a, &b, &r));
        std::vector<size_t> out(r.ids, r.ids + r.n_ids);
        pgi_locate_result_free(&r);
        return out;
    }
    std::vector<size_t> locate(range_type state, size_t first = NO_POSITION) const { return locate_encoded(state, first); }
    // decompressSA / decompressDA, src/r-index.cpp:1345-1365: SA value (seq_id * max_length + seq_offset) / sequence id of every row.
    std::vector<size_t> decompressSA() const {
        if (!ix_) throw std::runtime_error("FastLocate: no index loaded");
        if (!info_.n) return {};
        const uint64_t a = 0, b = info_.n - 1; pgi_locate_result r;
        pgi_check(pgi_locate_batch(ix_, 1, &a, &b, &r));
        std::vector<size_t> out(r.sa, r.sa + r.n_rows);
        pgi_locate_result_free(&r);
        return out;
    }
    std::vector<size_t> decompressDA() const { auto v = decompressSA(); for (auto &x : v) x = seqId(x); return v; }
    size_t seqId(size_t offset) const { return offset / info_.max_length; }     // r-index.hpp:429
    size_t seqOffset(size_t offset) const { return offset % info_.max_length; } // r-index.hpp:431

    pgi_index *handle() const { return ix_; }
    const std::string &file_bytes() const { return bytes_; }
    int device() const { return device_; }

private:
    int device_ = 0; pgi_index *ix_ = nullptr; pgi_index_info_t info_{}; std::string bytes_;
    void create() {
        if (ix_) { pgi_index_free(ix_); ix_ = nullptr; }
        pgi_check(pgi_index_create(bytes_.data(), bytes_.size(), nullptr, 0, 0, device_, &ix_));
        pgi_check(pgi_index_info(ix_, &info_));
        for (int i = 0; i < 256; ++i) sym_map[i] = info_.sym_map[i];
        C.assign(info_.C, info_.C + info_.sigma);
    }
    bi_interval extend(const bi_interval &b, size_t symbol, uint8_t dir) {
        if (!ix_) throw std::runtime_error("FastLocate: no index loaded");
        pgi_bi_interval in{b.forward, b.reverse, b.size}, out{0, 0, 0};
        const uint8_t sym = (uint8_t) symbol;
        pgi_check(pgi_extend_batch(ix_, 1, &in, &sym, &dir, &out));
        return bi_interval(out.forward, out.reverse, out.size);
    }
};

class TagArray {
public:
    TagArray() = default;
    explicit TagArray(int device) : device_(device) {}
    TagArray(const TagArray &) = delete;
    TagArray &operator=(const TagArray &) = delete;
    ~TagArray() { if (ix_) pgi_index_free(ix_); }

    void load_compressed_tags_compact(std::istream &in) { bytes_ = slurp(in); layout_ = PGI_TAGS_SDSL_COMPACT; create(); }
    void load_compressed_tags_sdsl(std::istream &in) { load_compressed_tags_compact(in); }
    void load_compressed_tags(std::istream &in) { bytes_ = slurp(in); layout_ = PGI_TAGS_BYTECODE_LEN; create(); }

    // Sorted unique encoded graph positions of the tag runs overlapping BWT rows [start, end]
    // (the reference's decode starts one run late when the first run index is a multiple of 10; reproduced).
    std::vector<uint64_t> query(size_t start, size_t end, size_t &number_of_runs) {
        if (!ix_) throw std::runtime_error("TagArray: no tags loaded");
        uint64_t s = start, e = end; pgi_result r;
        pgi_check(pgi_tag_query_batch(ix_, 1, &s, &e, &r));
        number_of_runs = (size_t) r.n_runs[0];
        std::vector<uint64_t> v(r.tags + r.tag_off[0], r.tags + r.tag_off[1]);
        pgi_result_free(&r);
        return v;
    }
    // tag_arrays.cpp:856-890 / :780-854: compute, then print "Number of unique positions: u" and the list.
    void query_compressed_compact(size_t start, size_t end, size_t &number_of_runs) { print(query(start, end, number_of_runs)); }
    void query_compressed(size_t start, size_t end, size_t &number_of_runs) { print(query(start, end, number_of_runs)); }

    pgi_index *handle() const { return ix_; }
    const std::string &file_bytes() const { return bytes_; }
    int layout() const { return layout_; }

private:
    int device_ = 0, layout_ = 0; pgi_index *ix_ = nullptr; std::string bytes_;
    void create() {
        if (ix_) { pgi_index_free(ix_); ix_ = nullptr; }
        pgi_check(pgi_index_create(nullptr, 0, bytes_.data(), bytes_.size(), layout_, device_, &ix_));
    }
    static void print(const std::vector<uint64_t> &v) {
        std::cout << "Number of unique positions: " << v.size() << std::endl;
        for (uint64_t x : v) std::cout << x << ", ";
        std::cout << std::endl;
    }
};

struct MEM {                                                                 // algorithm.hpp:644-651
    size_t start, end, bwt_start; int64_t size;
};

// algorithm.hpp:739-757. One read per call (one GPU launch): kept for source compatibility.
inline std::vector<MEM> find_all_mems(const std::string &pattern, size_t min_len, size_t min_occ, FastLocate &fmd_index) {
    if (!fmd_index.handle()) throw std::runtime_error("find_all_mems: no index loaded");
    const uint64_t off[2] = {0, (uint64_t) pattern.size()};
    pgi_result r;
    pgi_check(pgi_find_mems(fmd_index.handle(), pattern.data(), off, 1, min_len, min_occ, nullptr, &r));
    std::vector<MEM> out(r.n_mems);
    for (uint64_t i = 0; i < r.n_mems; ++i) out[i] = MEM{(size_t) r.mems[i].start, (size_t) r.mems[i].end, (size_t) r.mems[i].bwt_start, r.mems[i].size};
    pgi_result_free(&r);
    return out;
}

// The batched form: all reads of a batch in one pass. result[i] = MEMs of reads[i] in discovery order.
inline std::vector<std::vector<MEM>> find_all_mems_batch(const std::vector<std::string> &reads, size_t min_len, size_t min_occ, FastLocate &fmd_index) {
    if (!fmd_index.handle()) throw std::runtime_error("find_all_mems_batch: no index loaded");
    std::string buf; std::vector<uint64_t> off(reads.size() + 1, 0);
    for (size_t i = 0; i < reads.size(); ++i) { buf += reads[i]; off[i + 1] = buf.size(); }
    pgi_result r;
    pgi_check(pgi_find_mems(fmd_index.handle(), buf.data(), off.data(), reads.size(), min_len, min_occ, nullptr, &r));
    std::vector<std::vector<MEM>> out(reads.size());
    for (size_t i = 0; i < reads.size(); ++i)
        for (uint64_t m = r.mem_off[i]; m < r.mem_off[i + 1]; ++m)
            out[i].push_back(MEM{(size_t) r.mems[m].start, (size_t) r.mems[m].end, (size_t) r.mems[m].bwt_start, r.mems[m].size});
    pgi_result_free(&r);
    return out;
}

} // namespace panindexer
