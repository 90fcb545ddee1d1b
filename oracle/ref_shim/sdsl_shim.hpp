// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).
//
// A minimal stand-in for the parts of SDSL (vgteam fork) that the reference's OWN, UNMODIFIED sources
// src/r-index.cpp and src/tag_arrays.cpp call, so that those two files compile where they lie under
// /root/reference (oracle/Makefile -> oracle/_ref/libpgiref.so). The reference's dependencies (sdsl-lite, gbwt,
// gbwtgraph, libhandlegraph) are neither vendored nor installed here; this header restates only their
// PUBLISHED container semantics and on-disk layouts, nothing of the reference:
//   int_vector<W> / bit_vector     u64 size_in_bits, [u8 width iff W == 0], ceil(bits / 64) u64 words
//   sd_vector<> + sd_vector_builder Elias-Fano: u64 size, u8 wl, int_vector<0> low, bit_vector high, two
//                                  select_support_mcl blobs; rank_1 / select_1 / predecessor
//   int_vector_buffer<W>           append-only file writer producing the int_vector layout
//   structure_tree, util, bits, write_member / read_member, serialize / load / size_in_bytes, simple_sds::InvalidData
// Layouts are pinned by the shipped fixtures: the reference's own load() reads test_data/**/xy.ri and
// xy_bidirectional_compressed.tags through this shim to EOF, and its serialize() reproduces the files byte for byte
// (tests/test_ref_pin.py). Operations that are undefined in SDSL (select past the last one, operator[] past the end)
// are DEFINED here the way oracle/ defines them so that the comparison is total.
//
// None of this is linked, imported or executed by the product (pangenome-index_b200/).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <stdexcept>
#include <string>
#include <typeinfo>
#include <vector>

namespace sdsl {

typedef std::uint64_t shim_size_type;

struct structure_tree_node {};
struct structure_tree {
    static structure_tree_node *add_child(structure_tree_node *, const std::string &, const std::string &) { return nullptr; }
    static void add_size(structure_tree_node *, std::uint64_t) {}
};

namespace bits {
inline std::uint32_t hi(std::uint64_t x) { return x ? 63u - (std::uint32_t) __builtin_clzll(x) : 0u; }
inline std::uint32_t lo(std::uint64_t x) { return x ? (std::uint32_t) __builtin_ctzll(x) : 0u; }
inline std::uint32_t cnt(std::uint64_t x) { return (std::uint32_t) __builtin_popcountll(x); }
// bits needed to write x (0 for 0); widths of the shipped .ri files (13 for samples, ...) agree
inline std::uint8_t length(std::uint64_t x) { return x ? (std::uint8_t) (hi(x) + 1) : 0; }
} // namespace bits

template <class T> inline std::uint64_t write_member(const T &t, std::ostream &out, structure_tree_node * = nullptr, std::string = "") {
    out.write(reinterpret_cast<const char *>(&t), sizeof(T));
    return sizeof(T);
}
template <class T> inline void read_member(T &t, std::istream &in) { in.read(reinterpret_cast<char *>(&t), sizeof(T)); }

namespace simple_sds {
struct InvalidData : public std::runtime_error { using std::runtime_error::runtime_error; };
struct CannotOpenFile : public std::runtime_error { using std::runtime_error::runtime_error; };
} // namespace simple_sds

// ------------------------------------------------------------------------------------------ int_vector
template <std::uint8_t W> class int_vector;

template <std::uint8_t W> class int_vector_ref {
public:
    int_vector_ref(int_vector<W> *v, std::uint64_t i) : v_(v), i_(i) {}
    operator std::uint64_t() const { return v_->get(i_); }
    int_vector_ref &operator=(std::uint64_t x) { v_->set(i_, x); return *this; }
    int_vector_ref &operator=(const int_vector_ref &o) { v_->set(i_, (std::uint64_t) o); return *this; }
    int_vector_ref &operator+=(std::uint64_t x) { v_->set(i_, v_->get(i_) + x); return *this; }
    int_vector_ref &operator-=(std::uint64_t x) { v_->set(i_, v_->get(i_) - x); return *this; }
    int_vector_ref &operator++() { v_->set(i_, v_->get(i_) + 1); return *this; }
    std::uint64_t operator++(int) { const std::uint64_t o = v_->get(i_); v_->set(i_, o + 1); return o; }
    bool operator==(std::uint64_t x) const { return v_->get(i_) == x; }
private:
    int_vector<W> *v_; std::uint64_t i_;
};

template <std::uint8_t W> class int_vector {
public:
    typedef std::uint64_t size_type;
    typedef std::uint64_t value_type;
    typedef int_vector_ref<W> reference;
    class const_iterator {
    public:
        typedef std::random_access_iterator_tag iterator_category; typedef std::uint64_t value_type; typedef std::ptrdiff_t difference_type;
        typedef const std::uint64_t *pointer; typedef std::uint64_t reference;
        const_iterator(const int_vector *v = nullptr, std::uint64_t i = 0) : v_(v), i_(i) {}
        std::uint64_t operator*() const { return v_->get(i_); }
        const_iterator &operator++() { ++i_; return *this; }
        const_iterator operator++(int) { const_iterator t = *this; ++i_; return t; }
        const_iterator &operator--() { --i_; return *this; }
        const_iterator &operator+=(difference_type d) { i_ += d; return *this; }
        const_iterator operator+(difference_type d) const { return const_iterator(v_, i_ + d); }
        difference_type operator-(const const_iterator &o) const { return (difference_type) i_ - (difference_type) o.i_; }
        bool operator==(const const_iterator &o) const { return i_ == o.i_; }
        bool operator!=(const const_iterator &o) const { return i_ != o.i_; }
        bool operator<(const const_iterator &o) const { return i_ < o.i_; }
    private:
        const int_vector *v_; std::uint64_t i_;
    };
    // select / rank types only make sense for W == 1 (bit_vector); declared below
    struct select_1_type;
    struct rank_1_type;

    int_vector(size_type n = 0, value_type def = 0, std::uint8_t w = W ? W : 64) : width_(W ? W : w), size_(0) { resize(n); if (def) for (size_type i = 0; i < n; ++i) set(i, def); }
    std::uint8_t width() const { return width_; }
    void width(std::uint8_t w) { if (W == 0) { width_ = w; words_.assign(words_for(size_), 0); } }
    size_type size() const { return size_; }
    bool empty() const { return size_ == 0; }
    size_type bit_size() const { return size_ * width_; }
    size_type capacity() const { return (std::uint64_t) words_.size() * 64; }
    void resize(size_type n) { size_ = n; words_.resize(words_for(n), 0); }
    void bit_resize(size_type bits) { size_ = width_ ? bits / width_ : 0; words_.resize((std::size_t) ((bits + 63) >> 6) + 1, 0); }
    void clear() { size_ = 0; words_.clear(); }
    const std::uint64_t *data() const { return words_.data(); }
    std::uint64_t *data() { return words_.data(); }
    // reading past the end yields 0 (undefined in SDSL; defined like oracle/)
    value_type get(size_type i) const {
        if (i >= size_ || width_ == 0) return 0;
        const std::uint64_t b = i * width_, w = b >> 6, s = b & 63;
        std::uint64_t v = words_[w] >> s;
        if (s + width_ > 64) v |= words_[w + 1] << (64 - s);
        return width_ == 64 ? v : (v & ((1ULL << width_) - 1));
    }
    void set(size_type i, value_type v) {
        if (i >= size_ || width_ == 0) return;
        const std::uint64_t b = i * width_, w = b >> 6, s = b & 63;
        const std::uint64_t mask = width_ == 64 ? ~0ULL : ((1ULL << width_) - 1);
        v &= mask;
        words_[w] = (words_[w] & ~(mask << s)) | (v << s);
        if (s + width_ > 64) { const unsigned r = 64 - (unsigned) s; words_[w + 1] = (words_[w + 1] & ~(mask >> r)) | (v >> r); }
    }
    value_type operator[](size_type i) const { return get(i); }
    reference operator[](size_type i) { return reference(this, i); }
    const_iterator begin() const { return const_iterator(this, 0); }
    const_iterator end() const { return const_iterator(this, size_); }
    void swap(int_vector &o) { std::swap(width_, o.width_); std::swap(size_, o.size_); words_.swap(o.words_); }
    size_type serialize(std::ostream &out, structure_tree_node * = nullptr, std::string = "") const {
        const std::uint64_t bits = size_ * width_; size_type written = 0;
        written += write_member(bits, out);
        if (W == 0) written += write_member(width_, out);
        const std::size_t nw = (std::size_t) ((bits + 63) >> 6);
        if (nw) out.write(reinterpret_cast<const char *>(words_.data()), (std::streamsize) (nw * 8));
        return written + nw * 8;
    }
    void load(std::istream &in) {
        std::uint64_t bits = 0; read_member(bits, in);
        if (W == 0) read_member(width_, in);
        if (!in) throw simple_sds::InvalidData("int_vector: truncated input");
        size_ = width_ ? bits / width_ : 0;
        const std::size_t nw = (std::size_t) ((bits + 63) >> 6);
        words_.assign(nw + 1, 0);
        if (nw) in.read(reinterpret_cast<char *>(words_.data()), (std::streamsize) (nw * 8));
        if (!in) throw simple_sds::InvalidData("int_vector: truncated input");
    }
    bool operator==(const int_vector &o) const {
        if (size_ != o.size_ || width_ != o.width_) return false;
        for (size_type i = 0; i < size_; ++i) if (get(i) != o.get(i)) return false;
        return true;
    }
private:
    std::size_t words_for(size_type n) const { return (std::size_t) ((n * width_ + 63) >> 6) + 1; }   // + 1: two-word reads stay inside
    std::uint8_t width_; size_type size_; std::vector<std::uint64_t> words_;
};

typedef int_vector<1> bit_vector;

// bit_vector::select_1_type / rank_1_type: plain scans are enough (the reference only uses them on build paths)
template <std::uint8_t W> struct int_vector<W>::select_1_type {
    const int_vector<W> *v = nullptr;
    select_1_type() {}
    explicit select_1_type(const int_vector<W> *p) : v(p) {}
    void set_vector(const int_vector<W> *p) { v = p; }
    std::uint64_t operator()(std::uint64_t i) const {                    // position of the i-th (1-based) one
        std::uint64_t seen = 0;
        for (std::uint64_t p = 0; p < v->size(); ++p) if (v->get(p) && ++seen == i) return p;
        return v->size();
    }
    std::uint64_t serialize(std::ostream &, structure_tree_node * = nullptr, std::string = "") const { return 0; }
    void load(std::istream &, const int_vector<W> *p = nullptr) { v = p; }
};
template <std::uint8_t W> struct int_vector<W>::rank_1_type {
    const int_vector<W> *v = nullptr;
    rank_1_type() {}
    explicit rank_1_type(const int_vector<W> *p) : v(p) {}
    void set_vector(const int_vector<W> *p) { v = p; }
    std::uint64_t operator()(std::uint64_t i) const { std::uint64_t r = 0; for (std::uint64_t p = 0; p < i && p < v->size(); ++p) r += v->get(p); return r; }
};

namespace util {
template <class T> inline std::string class_name(const T &) { return typeid(T).name(); }
template <class T> inline void clear(T &x) { T y; x.swap(y); }
template <class V> inline void set_to_value(V &v, std::uint64_t k) { for (std::uint64_t i = 0; i < v.size(); ++i) v[i] = k; }
template <class V> inline void assign(V &a, V &b) { a.swap(b); }
template <class V> inline void bit_compress(V &) {}
} // namespace util

// ------------------------------------------------------------------------------------------ sd_vector
// select_support_mcl<b, 1> blob (layout only; content as SDSL writes it, so serialize() is byte-exact)
inline std::uint64_t shim_write_select_support(std::ostream &out, const std::vector<std::uint64_t> &bv, std::uint64_t nbits, int b) {
    const std::uint64_t SB = 4096;
    auto word = [&](std::uint64_t w) { std::uint64_t x = b ? bv[w] : ~bv[w]; const std::uint64_t rem = nbits - w * 64; if (rem < 64) x &= (1ULL << rem) - 1; return x; };
    std::uint64_t arg_cnt = 0;
    for (std::uint64_t w = 0; w * 64 < nbits; ++w) arg_cnt += bits::cnt(word(w));
    std::uint64_t written = write_member(arg_cnt, out);
    if (arg_cnt == 0) return written;
    const std::uint64_t capacity = ((nbits + 63) >> 6) << 6;
    const unsigned logn = bits::hi(capacity) + 1;
    const std::uint64_t logn4 = (std::uint64_t) logn * logn * logn * logn;
    const std::uint64_t sb = (arg_cnt + SB - 1) / SB;
    int_vector<0> superblock(sb, 0, (std::uint8_t) logn);
    std::vector<int_vector<0>> blocks((std::size_t) sb);
    std::vector<std::uint8_t> is_mini((std::size_t) sb, 0);
    bool any_long = false;
    std::vector<std::uint64_t> argpos(SB);
    std::uint64_t cnt = 0, sbi = 0;
    for (std::uint64_t w = 0; w * 64 < nbits; ++w) {
        std::uint64_t x = word(w);
        while (x) {
            const unsigned bit = bits::lo(x); x &= x - 1;
            argpos[cnt % SB] = w * 64 + bit; ++cnt;
            if (cnt % SB == 0 || cnt == arg_cnt) {
                const std::uint64_t lasti = (cnt - 1) % SB;
                superblock[sbi] = argpos[0];
                const std::uint64_t pos_diff = argpos[lasti] - argpos[0];
                if (pos_diff > logn4) {
                    any_long = true;
                    blocks[sbi] = int_vector<0>(SB, 0, (std::uint8_t) (bits::hi(argpos[lasti]) + 1));
                    for (std::uint64_t j = 0; j <= lasti; ++j) blocks[sbi][j] = argpos[j];
                } else {
                    is_mini[sbi] = 1;
                    blocks[sbi] = int_vector<0>(64, 0, (std::uint8_t) (bits::hi(pos_diff) + 1));
                    for (std::uint64_t j = 0; j <= lasti; j += 64) blocks[sbi][j / 64] = argpos[j] - argpos[0];
                }
                ++sbi;
            }
        }
    }
    written += superblock.serialize(out);
    bit_vector mol(any_long ? sb : 0, 0);
    if (any_long) for (std::uint64_t i = 0; i < sb; ++i) mol[i] = is_mini[i];
    written += mol.serialize(out);
    for (std::uint64_t i = 0; i < sb; ++i) written += blocks[i].serialize(out);
    return written;
}
inline void shim_skip_select_support(std::istream &in) {
    std::uint64_t cnt = 0; read_member(cnt, in);
    if (!cnt) return;
    int_vector<0> a; a.load(in);
    bit_vector b; b.load(in);
    for (std::uint64_t i = 0, sb = (cnt + 4095) / 4096; i < sb; ++i) a.load(in);
}

class sd_vector_builder {
public:
    sd_vector_builder() : n_(0), m_(0) {}
    sd_vector_builder(std::uint64_t n, std::uint64_t m) : n_(n), m_(m) { ones_.reserve((std::size_t) m); }
    void set(std::uint64_t i) { ones_.push_back(i); }
    void set_unsafe(std::uint64_t i) { ones_.push_back(i); }
    std::uint64_t size() const { return n_; }
    std::uint64_t capacity() const { return m_; }
    std::uint64_t items() const { return ones_.size(); }
    std::uint64_t tail() const { return ones_.empty() ? 0 : ones_.back() + 1; }
    std::uint64_t n_, m_; std::vector<std::uint64_t> ones_;
};

template <class A = void, class B = void, class C = void> class sd_vector {
public:
    typedef std::uint64_t size_type;
    struct one_iterator {
        const sd_vector *v; std::uint64_t j; mutable std::pair<std::uint64_t, std::uint64_t> cur;
        const std::pair<std::uint64_t, std::uint64_t> *operator->() const { cur = {j, j < v->ones_.size() ? v->ones_[(std::size_t) j] : v->size_}; return &cur; }
        std::pair<std::uint64_t, std::uint64_t> operator*() const { return *operator->(); }
        one_iterator &operator++() { ++j; return *this; }
        bool operator==(const one_iterator &o) const { return j == o.j; }
        bool operator!=(const one_iterator &o) const { return j != o.j; }
    };
    struct rank_1_type {
        const sd_vector *v = nullptr;
        rank_1_type() {}
        explicit rank_1_type(const sd_vector *p) : v(p) {}
        void set_vector(const sd_vector *p) { v = p; }
        std::uint64_t operator()(std::uint64_t i) const { return v->rank1(i); }
        std::uint64_t rank(std::uint64_t i) const { return v->rank1(i); }
    };
    struct select_1_type {
        const sd_vector *v = nullptr;
        select_1_type() {}
        explicit select_1_type(const sd_vector *p) : v(p) {}
        void set_vector(const sd_vector *p) { v = p; }
        // i is 1-based. Past the last one (undefined in SDSL): a position far behind every array, so that whatever is read
        // there reads as zero -- the definition oracle/ and the CUDA path use ("defined UB")
        std::uint64_t operator()(std::uint64_t i) const { return (i >= 1 && i <= v->ones_.size()) ? v->ones_[(std::size_t) (i - 1)] : (1ULL << 62); }
        std::uint64_t select(std::uint64_t i) const { return (*this)(i); }
    };
    typedef rank_1_type rank_0_type; typedef select_1_type select_0_type;

    sd_vector() : size_(0), m_cap_(0) {}
    explicit sd_vector(const sd_vector_builder &b) : size_(b.n_), m_cap_(b.m_), ones_(b.ones_) { build_dir(); }
    sd_vector(sd_vector_builder &b) : size_(b.n_), m_cap_(b.m_), ones_(b.ones_) { build_dir(); }
    explicit sd_vector(const bit_vector &bv) : size_(bv.size()), m_cap_(0) { for (std::uint64_t i = 0; i < bv.size(); ++i) if (bv[i]) ones_.push_back(i); m_cap_ = ones_.size(); build_dir(); }
    size_type size() const { return size_; }
    size_type ones() const { return ones_.size(); }
    bool empty() const { return size_ == 0; }
    bool operator[](size_type i) const { return std::binary_search(ones_.begin(), ones_.end(), i); }
    void swap(sd_vector &o) { std::swap(size_, o.size_); std::swap(m_cap_, o.m_cap_); ones_.swap(o.ones_); dir_.swap(o.dir_); std::swap(dshift_, o.dshift_); }
    // #ones in [0, i); i beyond the universe counts everything (undefined in SDSL)
    std::uint64_t rank1(std::uint64_t i) const {
        if (i >= size_) return ones_.size();
        const std::uint64_t b = i >> dshift_;
        return (std::uint64_t) (std::lower_bound(ones_.begin() + (std::ptrdiff_t) dir_[(std::size_t) b], ones_.begin() + (std::ptrdiff_t) dir_[(std::size_t) b + 1], i) - ones_.begin());
    }
    one_iterator one_begin() const { return one_iterator{this, 0, {}}; }
    one_iterator one_end() const { return one_iterator{this, ones_.size(), {}}; }
    // last one at a position <= i (vgteam fork); one_end() when there is none
    one_iterator predecessor(std::uint64_t i) const {
        const std::uint64_t r = i >= size_ ? ones_.size() : rank1(i + 1);
        return r ? one_iterator{this, r - 1, {}} : one_end();
    }
    one_iterator successor(std::uint64_t i) const { return one_iterator{this, rank1(i), {}}; }
    // Elias-Fano parameters of sdsl::sd_vector_builder (simple-sds compatible): wl = round(max(log2(n ln 2 / m), 1))
    size_type serialize(std::ostream &out, structure_tree_node * = nullptr, std::string = "") const {
        const std::uint64_t m = std::max<std::uint64_t>(m_cap_, ones_.size()), n = size_;
        unsigned wl = 1;
        if (m > 0 && m <= n) { const double ideal = std::log2(((double) n * std::log(2.0)) / (double) m); wl = (unsigned) std::llround(std::max(ideal, 1.0)); }
        const std::uint64_t buckets = (n >> wl) + ((n & ((1ULL << wl) - 1)) ? 1 : 0), hbits = m + buckets;
        int_vector<0> low(m, 0, (std::uint8_t) wl);
        std::vector<std::uint64_t> high((std::size_t) ((hbits + 63) >> 6) + 1, 0);
        for (std::uint64_t j = 0; j < ones_.size(); ++j) {
            low[j] = ones_[(std::size_t) j] & ((1ULL << wl) - 1);
            const std::uint64_t hp = (ones_[(std::size_t) j] >> wl) + j; high[(std::size_t) (hp >> 6)] |= 1ULL << (hp & 63);
        }
        size_type written = write_member(n, out) + write_member((std::uint8_t) wl, out);
        written += low.serialize(out);
        written += write_member(hbits, out);
        const std::size_t nw = (std::size_t) ((hbits + 63) >> 6);
        if (nw) out.write(reinterpret_cast<const char *>(high.data()), (std::streamsize) (nw * 8));
        written += nw * 8;
        written += shim_write_select_support(out, high, hbits, 1);
        written += shim_write_select_support(out, high, hbits, 0);
        return written;
    }
    void load(std::istream &in) {
        std::uint8_t wl = 0; read_member(size_, in); read_member(wl, in);
        int_vector<0> low; low.load(in);
        bit_vector high; high.load(in);
        shim_skip_select_support(in); shim_skip_select_support(in);
        if (!in) throw simple_sds::InvalidData("sd_vector: truncated input");
        ones_.clear();
        std::uint64_t j = 0;
        const std::uint64_t *hw = high.data();
        for (std::uint64_t w = 0; w * 64 < high.size(); ++w) {
            std::uint64_t x = hw[w];
            while (x) {
                const unsigned b = bits::lo(x); x &= x - 1;
                const std::uint64_t pbit = w * 64 + b;
                if (pbit >= high.size()) break;
                ones_.push_back(((pbit - j) << wl) | (wl ? low.get(j) : 0)); ++j;
            }
        }
        m_cap_ = ones_.size();
        build_dir();
    }
private:
    void build_dir() {
        const std::uint64_t m = ones_.size();
        dshift_ = 0;
        if (m == 0) { dshift_ = 63; dir_.assign(3, 0); return; }
        while (dshift_ < 62 && ((size_ >> dshift_) > m / 2 + 1)) ++dshift_;
        const std::uint64_t nb = (size_ >> dshift_) + 2;
        dir_.assign((std::size_t) nb + 1, 0);
        std::uint64_t j = 0;
        for (std::uint64_t b = 0; b <= nb; ++b) { const std::uint64_t lim = b << dshift_; while (j < m && ones_[(std::size_t) j] < lim) ++j; dir_[(std::size_t) b] = j; }
    }
    std::uint64_t size_, m_cap_; std::vector<std::uint64_t> ones_;
    unsigned dshift_ = 63; std::vector<std::uint64_t> dir_;
};

// ------------------------------------------------------------------------------------------ free functions
template <class T> inline std::uint64_t serialize(const T &t, std::ostream &out, structure_tree_node *v = nullptr, std::string name = "") { return t.serialize(out, v, name); }
template <class T> inline void load(T &t, std::istream &in) { t.load(in); }
struct shim_null_buf : public std::streambuf { int overflow(int c) override { return c; } std::streamsize xsputn(const char *, std::streamsize n) override { return n; } };
template <class T> inline std::uint64_t size_in_bytes(const T &t) { shim_null_buf nb; std::ostream os(&nb); return t.serialize(os, nullptr, ""); }
template <class T> inline bool store_to_file(const T &t, const std::string &file) { std::ofstream out(file, std::ios::binary); if (!out) return false; t.serialize(out); return (bool) out; }
template <class T> inline bool load_from_file(T &t, const std::string &file) { std::ifstream in(file, std::ios::binary); if (!in) return false; t.load(in); return true; }

// ------------------------------------------------------------------------------------------ int_vector_buffer
// Append-only writer: collects the values and writes the int_vector layout when closed / destroyed.
template <std::uint8_t W> class int_vector_buffer {
public:
    int_vector_buffer() {}
    int_vector_buffer(const std::string &file, std::ios::openmode mode = std::ios::in, std::uint64_t = 1024 * 1024, std::uint8_t width = W ? W : 64, bool = false)
        : file_(file), v_(0, 0, W ? W : width), open_(true) {
        if (!(mode & std::ios::trunc) && (mode & std::ios::in)) { std::ifstream in(file, std::ios::binary); if (in) v_.load(in); }
    }
    int_vector_buffer(int_vector_buffer &&o) { *this = std::move(o); }
    int_vector_buffer &operator=(int_vector_buffer &&o) { close(); file_ = o.file_; v_.swap(o.v_); n_ = o.n_; open_ = o.open_; o.open_ = false; return *this; }
    ~int_vector_buffer() { close(); }
    void push_back(std::uint64_t x) { if (n_ >= v_.size()) v_.resize(std::max<std::uint64_t>(1024, 2 * v_.size())); v_.set(n_++, x); }
    std::uint64_t size() const { return n_; }
    std::uint8_t width() const { return v_.width(); }
    std::uint64_t operator[](std::uint64_t i) const { return v_.get(i); }
    bool is_open() const { return open_; }
    void close(bool remove_file = false) {
        if (!open_) return;
        open_ = false;
        if (remove_file) { std::remove(file_.c_str()); return; }
        v_.resize(n_);
        std::ofstream out(file_, std::ios::binary | std::ios::trunc); v_.serialize(out);
    }
private:
    std::string file_; int_vector<W> v_; std::uint64_t n_ = 0; bool open_ = false;
};

// ------------------------------------------------------------------------------------------ wt_huff (deps/grlBWT/scripts/fm_index.h only)
template <class A = void> class wt_huff {
public:
    std::vector<std::uint8_t> text; std::vector<std::vector<std::uint64_t>> occ;
    std::uint64_t size() const { return text.size(); }
    std::pair<std::uint64_t, std::uint8_t> inverse_select(std::uint64_t i) const {
        const std::uint8_t c = text[(std::size_t) i]; std::uint64_t r = 0;
        for (std::uint64_t p = 0; p < i; ++p) r += text[(std::size_t) p] == c;
        return {r, c};
    }
};
template <class WT> inline void construct(WT &wt, const std::string &file, std::uint8_t) {
    std::ifstream in(file, std::ios::binary); wt.text.assign(std::istreambuf_iterator<char>(in), std::istreambuf_iterator<char>());
}

} // namespace sdsl
