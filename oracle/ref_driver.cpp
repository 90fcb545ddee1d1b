// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_formats.hpp header).
//
// C ABI over the REFERENCE'S OWN CODE, for pinning oracle/ (the restatement) to what the reference computes.
// oracle/Makefile builds oracle/_ref/libpgiref.so from
//   * /root/reference/src/r-index.cpp and /root/reference/src/tag_arrays.cpp, UNMODIFIED, compiled where they lie,
//     against oracle/ref_shim/ (stand-ins for the un-vendored SDSL / GBWT / GBWTGraph / libhandlegraph headers);
//   * three passages cut out of reference files BY LINE RANGE at build time into oracle/_ref/*.inc, each checked
//     against a SHA-256 recorded in the Makefile, and #included below — never copied into this repository:
//       include/pangenome_index/algorithm.hpp:644-757   struct MEM, find_mems_function, find_all_mems
//       src/find_mems.cpp:93-139                        the per-read loop of main() (stdout grammar, tag query per MEM)
//       src/query_tags.cpp:87-109                       the per-read loop of the query_tags main()
//     (algorithm.hpp as a whole needs gbwtgraph::GBZ and friends for its index-construction half and does not compile
//     against the shim; the MEM finder itself only needs FastLocate.)
//   * this file: glue only (buffers <-> std::istream, std::cout capture, OpenMP over independent reads).
// The exported names and struct layouts are those of pgi_oracle.cpp, so tests/orc.py drives either library.
// Outputs that the reference only PRINTS (tag lists) are recovered by capturing std::cout around the call.
#include <cstdlib>
#include <cstring>
#include <memory>
#include <sstream>
#include <streambuf>
#include <string>
#include <vector>
#include <omp.h>

#define private public          // rank_at_cached_encoded, LF, psi are private in the reference; layout is unaffected
#include "pangenome_index/r-index.hpp"
#undef private
#include "pangenome_index/tag_arrays.hpp"

namespace gbwt { size_type Verbosity::level = Verbosity::SILENT; }

namespace panindexer {
#include "_ref/ref_algorithm_644_757.inc"
}

using namespace panindexer;

#define TIME 0
// src/find_mems.cpp:93-139 with the names it uses bound to parameters.
template <class TagQuery>
static void ref_find_mems_loop_generic(std::istream &reads, FastLocate &r_index, size_t mem_length, size_t min_occ, TagQuery &&query) {
    std::string read; int i = 0;
    while (std::getline(reads, read)) {
        if (read.empty()) continue;
        i++;
        auto mems = find_all_mems(read, mem_length, min_occ, r_index);
        std::cout << "Seq: " << i << std::endl;
        for (const auto &mem : mems) {
            std::cout << "MEM START: " << mem.start << ", MEM END: " << mem.end << " BWT START: " << mem.bwt_start << " SIZE: " << mem.size << std::endl;
            size_t tag_nums = 0;
            query(mem.bwt_start, mem.bwt_start + mem.size - 1, tag_nums);
        }
        std::cout << std::endl;
    }
}
static void ref_find_mems_loop(std::ifstream &reads, FastLocate &r_index, TagArray &tag_array, size_t mem_length, size_t min_occ) {
    using namespace std;
#include "_ref/ref_find_mems_93_139.inc"
}
static int ref_query_tags_loop(std::vector<std::string> &reads, FastLocate &r_index, TagArray &tag_array) {
#include "_ref/ref_query_tags_87_109.inc"
    return 0;
}

struct orc_index {
    FastLocate ri; TagArray tags; bool has_tags = false, has_ri = false; int layout = 0;
};
struct orc_result {
    uint64_t n_reads, n_mems, n_tags;
    uint64_t *mem_off; MEM *mems; uint64_t *tag_off; uint64_t *tags; uint64_t *n_runs;
    uint64_t counters[8];
};

static thread_local std::string g_err;

namespace {
struct MemBuf : std::streambuf {
    MemBuf(const uint8_t *p, size_t n) { char *b = const_cast<char *>(reinterpret_cast<const char *>(p)); setg(b, b, b + n); }
    pos_type seekoff(off_type off, std::ios_base::seekdir dir, std::ios_base::openmode) override {
        char *t = dir == std::ios_base::beg ? eback() + off : dir == std::ios_base::cur ? gptr() + off : egptr() + off;
        if (t < eback() || t > egptr()) return pos_type(off_type(-1));
        setg(eback(), t, egptr()); return pos_type(t - eback());
    }
    pos_type seekpos(pos_type pos, std::ios_base::openmode m) override { return seekoff(off_type(pos), std::ios_base::beg, m); }
    size_t consumed() const { return (size_t) (gptr() - eback()); }
};
// Everything the reference writes to std::cout while alive; std::cerr muted (find_all_mems logs one line per read).
struct Capture {
    std::ostringstream out; std::streambuf *old_out, *old_err; std::ios_base::iostate err_state;
    Capture() { old_out = std::cout.rdbuf(out.rdbuf()); err_state = std::cerr.rdstate(); old_err = std::cerr.rdbuf(); std::cerr.setstate(std::ios_base::badbit); }
    ~Capture() { std::cout.rdbuf(old_out); std::cerr.clear(err_state); std::cerr.rdbuf(old_err); }
};
struct Mute {
    std::ios_base::iostate o, e;
    Mute() { o = std::cout.rdstate(); e = std::cerr.rdstate(); std::cout.setstate(std::ios_base::badbit); std::cerr.setstate(std::ios_base::badbit); }
    ~Mute() { std::cout.clear(o); std::cerr.clear(e); }
};
// "Number of unique positions: <u>\n<p1>, <p2>, ..., \n" (src/tag_arrays.cpp:848-853 / :885-889) -> values
void parse_tag_text(const std::string &s, std::vector<uint64_t> &v) {
    v.clear();
    const size_t nl = s.find('\n');
    if (nl == std::string::npos) throw std::runtime_error("reference tag query printed no first line");
    const char *p = s.c_str() + nl + 1;
    while (*p && *p != '\n') {
        char *e = nullptr; const unsigned long long x = std::strtoull(p, &e, 10);
        if (e == p) break;
        v.push_back(x); p = e;
        while (*p == ',' || *p == ' ') ++p;
    }
    const unsigned long long claimed = std::strtoull(s.c_str() + std::strlen("Number of unique positions: "), nullptr, 10);
    if (claimed != v.size()) throw std::runtime_error("reference tag query: count line disagrees with the values printed");
}
uint64_t ref_tag_query(orc_index *ix, uint64_t start, uint64_t end, std::vector<uint64_t> &v) {
    size_t nr = 0; std::string text;
    { Capture c; if (ix->layout == 0) ix->tags.query_compressed_compact(start, end, nr); else ix->tags.query_compressed(start, end, nr); text = c.out.str(); }
    parse_tag_text(text, v);
    return nr;
}
char *dup_out(const std::string &s, uint64_t *len) {
    char *buf = (char *) std::malloc(s.size() + 1); std::memcpy(buf, s.data(), s.size()); buf[s.size()] = 0; *len = s.size(); return buf;
}
} // namespace

extern "C" {

const char *orc_last_error() { return g_err.c_str(); }
int ref_is_reference() { return 1; }

orc_index *orc_index_load(const uint8_t *ri, uint64_t ri_len, const uint8_t *tg, uint64_t tg_len, int tag_layout,
                          uint64_t *ri_consumed, uint64_t *tg_consumed) {
    try {
        std::unique_ptr<orc_index> ix(new orc_index());
        Mute m;
        if (ri && ri_len) {
            MemBuf b(ri, ri_len); std::istream in(&b);
            ix->ri.load_encoded(in);                                   // src/find_mems.cpp:31 (falls back to load() on a legacy file)
            if (!in) throw std::runtime_error("reference: truncated r-index");
            if (ri_consumed) *ri_consumed = b.consumed();
            ix->has_ri = true;
        }
        if (tg && tg_len) {
            MemBuf b(tg, tg_len); std::istream in(&b);
            if (tag_layout == 0) ix->tags.load_compressed_tags_compact(in);   // src/find_mems.cpp:79
            else if (tag_layout == 1) ix->tags.load_compressed_tags(in);     // src/find_mems_memory_mapped.cpp:93
            else throw std::runtime_error("unknown tag layout");
            if (!in) throw std::runtime_error("reference: truncated tags");
            if (tg_consumed) *tg_consumed = b.consumed();
            ix->has_tags = true; ix->layout = tag_layout;
        }
        return ix.release();
    } catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}
void orc_index_free(orc_index *ix) { delete ix; }

void orc_index_info(const orc_index *ix, uint64_t *info, uint64_t *C_out, uint8_t *sym_map_out) {
    const FastLocate &r = ix->ri;
    for (int i = 0; i < 16; ++i) info[i] = 0;
    info[0] = r.sequence_size; info[1] = r.C.size(); info[2] = r.is_encoded(); info[3] = r.encoded_has_N;
    info[4] = r.blocks_start_pos.ones(); info[5] = r.header.max_length; info[6] = r.header.flags;
    info[7] = r.samples.size(); info[8] = r.last.ones(); info[9] = r.last_to_run.size(); info[14] = r.samples.width();
    if (C_out) for (size_t i = 0; i < r.C.size(); ++i) C_out[i] = r.C[i];
    if (sym_map_out) for (size_t i = 0; i < 256 && i < r.sym_map.size(); ++i) sym_map_out[i] = (uint8_t) r.sym_map[i];
}

// FastLocate::backward_extend_encoded / forward_extend_encoded (src/r-index.cpp:713-764) themselves.
void orc_extend_batch(orc_index *ix, uint64_t n, const uint64_t *in3, const uint8_t *sym, const uint8_t *dir, uint64_t *out3) {
    for (uint64_t i = 0; i < n; ++i) {
        FastLocate::bi_interval b; b.forward = in3[3 * i]; b.reverse = in3[3 * i + 1]; b.size = (int64_t) in3[3 * i + 2];
        FastLocate::bi_interval r = dir[i] ? ix->ri.forward_extend_encoded(b, sym[i]) : ix->ri.backward_extend_encoded(b, sym[i]);
        out3[3 * i] = r.forward; out3[3 * i + 1] = r.reverse; out3[3 * i + 2] = (uint64_t) r.size;
    }
}
// FastLocate::rank_at_cached_encoded (src/r-index.cpp:619-641).
void orc_rank_batch(orc_index *ix, uint64_t n, const uint64_t *pos, uint64_t *out8) {
    for (uint64_t i = 0; i < n; ++i) {
        std::vector<size_t> r = ix->ri.rank_at_cached_encoded(pos[i]);
        for (size_t j = 0; j < 8; ++j) out8[8 * i + j] = j < r.size() ? r[j] : 0;
    }
}
uint64_t orc_tag_query(orc_index *ix, uint64_t start, uint64_t end, uint64_t *out, uint64_t cap, uint64_t *n_out) {
    try {
        std::vector<uint64_t> v; const uint64_t nr = ref_tag_query(ix, start, end, v);
        *n_out = v.size();
        for (size_t i = 0; i < v.size() && i < cap; ++i) out[i] = v[i];
        return nr;
    } catch (const std::exception &e) { g_err = e.what(); *n_out = 0; return ~0ull; }
}

// find_all_mems (algorithm.hpp:739-757) per read, in parallel over reads (the reference's own loop is serial; the MEM
// finder only reads the index). Tag queries print to std::cout, so they run serially under a capture afterwards.
int orc_find_mems(orc_index *ix, const char *reads, const uint64_t *read_off, uint64_t n_reads,
                  uint64_t min_len, uint64_t min_occ, int with_tags, int threads, int /*count_work*/, orc_result *res) {
    try {
        if (with_tags && !ix->has_tags) throw std::runtime_error("reference: tags not loaded");
        std::memset(res, 0, sizeof(*res));
        if (threads <= 0) threads = omp_get_max_threads();
        std::vector<std::vector<MEM>> per(n_reads);
        {
            Mute m;
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
            for (int64_t r = 0; r < (int64_t) n_reads; ++r) {
                const std::string pattern(reads + read_off[r], reads + read_off[r + 1]);
                per[r] = find_all_mems(pattern, min_len, min_occ, ix->ri);
            }
        }
        uint64_t nm = 0; for (auto &v : per) nm += v.size();
        res->n_reads = n_reads; res->n_mems = nm;
        res->mem_off = (uint64_t *) std::malloc((n_reads + 1) * 8); res->mems = (MEM *) std::malloc((nm + 1) * sizeof(MEM));
        res->tag_off = (uint64_t *) std::malloc((nm + 1) * 8); res->n_runs = (uint64_t *) std::malloc((nm + 1) * 8);
        std::vector<uint64_t> tags, tv;
        uint64_t mi = 0;
        for (uint64_t r = 0; r < n_reads; ++r) {
            res->mem_off[r] = mi;
            for (const MEM &m : per[r]) {
                res->mems[mi] = m; res->tag_off[mi] = tags.size(); res->n_runs[mi] = 0;
                if (with_tags) { res->n_runs[mi] = ref_tag_query(ix, m.bwt_start, m.bwt_start + m.size - 1, tv); tags.insert(tags.end(), tv.begin(), tv.end()); }
                ++mi;
            }
        }
        res->mem_off[n_reads] = nm; res->tag_off[nm] = tags.size(); res->n_tags = tags.size();
        res->tags = (uint64_t *) std::malloc((tags.size() + 1) * 8);
        if (!tags.empty()) std::memcpy(res->tags, tags.data(), tags.size() * 8);
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

// The timed form (bench.py --impl reference, cpu_baseline): MEM finding AND the tag query of every MEM inside the parallel
// loop, exactly the work of src/find_mems.cpp:95-139 per read, with the two streams muted (a failed stream skips the
// formatting, which only favours the reference). Returns seconds; *n_mems_out = MEMs found.
double ref_time_find_mems(orc_index *ix, const char *reads, const uint64_t *read_off, uint64_t n_reads,
                          uint64_t min_len, uint64_t min_occ, int with_tags, int threads, uint64_t *n_mems_out) {
    if (threads <= 0) threads = omp_get_max_threads();
    uint64_t total = 0;
    Mute m;
    const double t0 = omp_get_wtime();
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads) reduction(+ : total)
    for (int64_t r = 0; r < (int64_t) n_reads; ++r) {
        const std::string pattern(reads + read_off[r], reads + read_off[r + 1]);
        auto mems = find_all_mems(pattern, min_len, min_occ, ix->ri);
        total += mems.size();
        if (with_tags) for (const auto &mem : mems) {
            size_t tag_nums = 0;
            if (ix->layout == 0) ix->tags.query_compressed_compact(mem.bwt_start, mem.bwt_start + mem.size - 1, tag_nums);
            else ix->tags.query_compressed(mem.bwt_start, mem.bwt_start + mem.size - 1, tag_nums);
        }
    }
    const double dt = omp_get_wtime() - t0;
    if (n_mems_out) *n_mems_out = total;
    return dt;
}

// stdout of the reference's find_mems main loop over newline-separated reads. verbatim != 0 and tag layout 0: the loop
// cut out of src/find_mems.cpp:93-139 itself (it reads a std::ifstream, so the reads go through a temporary file);
// otherwise the same loop restated above with query_compressed for the ByteCode layout (find_mems_memory_mapped.cpp:93).
char *ref_find_mems_text(orc_index *ix, const char *reads_text, uint64_t text_len, uint64_t min_len, uint64_t min_occ, int verbatim, uint64_t *len) {
    try {
        std::string text;
        if (verbatim && ix->layout == 0 && ix->has_tags) {
            char path[] = "/tmp/pgiref_reads_XXXXXX"; const int fd = mkstemp(path);
            if (fd < 0) throw std::runtime_error("mkstemp failed");
            FILE *f = fdopen(fd, "wb"); fwrite(reads_text, 1, text_len, f); fclose(f);
            std::ifstream in(path);
            { Capture c; ref_find_mems_loop(in, ix->ri, ix->tags, min_len, min_occ); text = c.out.str(); }
            std::remove(path);
        } else {
            std::istringstream in(std::string(reads_text, text_len));
            Capture c;
            if (!ix->has_tags) ref_find_mems_loop_generic(in, ix->ri, min_len, min_occ, [](size_t, size_t, size_t &) {});
            else if (ix->layout == 0) ref_find_mems_loop_generic(in, ix->ri, min_len, min_occ, [&](size_t a, size_t b, size_t &n) { ix->tags.query_compressed_compact(a, b, n); });
            else ref_find_mems_loop_generic(in, ix->ri, min_len, min_occ, [&](size_t a, size_t b, size_t &n) { ix->tags.query_compressed(a, b, n); });
            text = c.out.str();
        }
        return dup_out(text, len);
    } catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}
// stdout of the loop cut out of src/query_tags.cpp:87-109 (sdsl-compact tags only, as that main loads them).
char *ref_query_tags_text(orc_index *ix, const char *reads_text, uint64_t text_len, uint64_t *len) {
    try {
        std::vector<std::string> reads; std::istringstream in(std::string(reads_text, text_len)); std::string line;
        while (std::getline(in, line)) if (!line.empty()) reads.push_back(line);
        std::string text;
        { Capture c; ref_query_tags_loop(reads, ix->ri, ix->tags); text = c.out.str(); }
        return dup_out(text, len);
    } catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}

// FastLocate::locate_encoded / locate (src/r-index.cpp:1256-1343) and decompressSA (:1345-1356).
uint64_t orc_locate_batch(orc_index *ix, uint64_t n, const uint64_t *first, const uint64_t *second,
                          uint64_t *out_off, uint64_t *out, uint64_t cap, uint64_t *sa_out, uint64_t sa_cap) {
    uint64_t tot = 0; (void) sa_out; (void) sa_cap;
    for (uint64_t i = 0; i < n; ++i) {
        std::vector<gbwt::size_type> v = ix->ri.is_encoded() ? ix->ri.locate_encoded({first[i], second[i]}) : ix->ri.locate({first[i], second[i]});
        out_off[i] = tot;
        for (uint64_t x : v) { if (tot < cap) out[tot] = x; ++tot; }
    }
    out_off[n] = tot;
    return tot;
}
void orc_decompress_sa(orc_index *ix, uint64_t *out) {
    std::vector<gbwt::size_type> v = ix->ri.decompressSA();
    for (size_t i = 0; i < v.size(); ++i) out[i] = v[i];
}

// FastLocate(std::string rlbwt_path) (src/r-index.cpp:778-1139) + serialize_encoded / serialize (:266-376): the bytes
// of the .ri file the reference's build_rindex writes (src/build_rindex.cpp:11-20).
char *ref_build_rindex(const char *rlbwt_path, int encoded, uint64_t *len) {
    try {
        std::ostringstream os;
        { Mute m; FastLocate idx{std::string(rlbwt_path)}; if (encoded) idx.serialize_encoded(os); else idx.serialize(os); }
        return dup_out(os.str(), len);
    } catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}

void orc_result_free(orc_result *r) {
    std::free(r->mem_off); std::free(r->mems); std::free(r->tag_off); std::free(r->tags); std::free(r->n_runs);
    std::memset(r, 0, sizeof(*r));
}
void orc_free(void *p) { std::free(p); }
int orc_max_threads() { return omp_get_max_threads(); }

} // extern "C"
