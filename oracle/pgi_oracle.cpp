// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_formats.hpp header).
//
// C ABI over the CPU restatement so tests/ and bench.py's cpu_baseline leg can drive it
// through ctypes. The batch driver parallelises independent reads with OpenMP; the
// reference's own find_mems loop is serial (src/find_mems.cpp:94-139), so threads>1 is
// *our* parallelisation of the reference algorithm and is reported as such.
#include "orc_mems.hpp"
#include <cstdlib>
#include <cstdio>
#include <memory>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

struct orc_index {
    FastLocate ri;
    TagArray tags;
    bool has_tags = false;
};

struct orc_result {
    uint64_t n_reads, n_mems, n_tags;
    uint64_t *mem_off;      // n_reads + 1
    MEM *mems;              // n_mems   (start, end, bwt_start, size)
    uint64_t *tag_off;      // n_mems + 1
    uint64_t *tags;         // n_tags
    uint64_t *n_runs;       // n_mems
    uint64_t counters[8];   // E, D, M, tag_runs, tag_skipped, tag_item_sectors, 0, 0
};

static thread_local std::string g_err;

extern "C" {

const char *orc_last_error() { return g_err.c_str(); }

// tag_layout: 0 sdsl-compact (load_compressed_tags_compact), 1 ByteCode+len9 (load_compressed_tags); tag bytes may be NULL.
orc_index *orc_index_load(const uint8_t *ri, uint64_t ri_len, const uint8_t *tg, uint64_t tg_len, int tag_layout,
                          uint64_t *ri_consumed, uint64_t *tg_consumed) {
    try {
        std::unique_ptr<orc_index> ix(new orc_index());
        Cursor c(ri, ri_len);
        ix->ri.load_encoded(c);
        if (ri_consumed) *ri_consumed = c.off;
        if (tg && tg_len) {
            Cursor t(tg, tg_len);
            if (tag_layout == 0) ix->tags.load_compressed_tags_compact(t);
            else if (tag_layout == 1) ix->tags.load_compressed_tags(t);
            else throw std::runtime_error("oracle: unknown tag layout");
            if (tg_consumed) *tg_consumed = t.off;
            ix->has_tags = true;
        }
        return ix.release();
    } catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}
void orc_index_free(orc_index *ix) { delete ix; }

// info[0..15]: n, sigma, is_encoded, hasN, n_blocks(ones in blocks_start_pos), max_length, flags,
//              samples.size, last.ones, last_to_run.size, tag items / stream bytes, tag starts ones, bwt_intervals ones,
//              bwt_intervals.size, samples.width, tag item width
void orc_index_info(const orc_index *ix, uint64_t *info, uint64_t *C_out, uint8_t *sym_map_out) {
    const FastLocate &r = ix->ri;
    info[0] = r.sequence_size; info[1] = r.C.size(); info[2] = r.is_encoded(); info[3] = r.encoded_has_N;
    info[4] = r.blocks_start_pos.ones.size(); info[5] = r.max_length; info[6] = r.flags;
    info[7] = r.samples.size(); info[8] = r.last.ones.size(); info[9] = r.last_to_run.size();
    info[10] = ix->tags.layout == 0 ? ix->tags.encoded_runs_iv.size() : ix->tags.encoded_runs.size();
    info[11] = ix->tags.encoded_runs_starts_sd.ones.size(); info[12] = ix->tags.bwt_intervals.ones.size();
    info[13] = ix->tags.bwt_intervals.size; info[14] = r.samples.width; info[15] = ix->tags.encoded_runs_iv.width;
    if (C_out) for (size_t i = 0; i < r.C.size(); ++i) C_out[i] = r.C[i];
    if (sym_map_out) std::memcpy(sym_map_out, r.sym_map, 256);
}

// Batched FMD extension: io = n x {forward, reverse, size}; dir 0 backward, 1 forward.
void orc_extend_batch(const orc_index *ix, uint64_t n, const uint64_t *in3, const uint8_t *sym, const uint8_t *dir, uint64_t *out3) {
    for (uint64_t i = 0; i < n; ++i) {
        bi_interval b{in3[3 * i], in3[3 * i + 1], (int64_t) in3[3 * i + 2]};
        bi_interval r = dir[i] ? ix->ri.forward_extend_encoded(b, sym[i]) : ix->ri.backward_extend_encoded(b, sym[i]);
        out3[3 * i] = r.forward; out3[3 * i + 1] = r.reverse; out3[3 * i + 2] = (uint64_t) r.size;
    }
}

// Rank vectors R(pos)[0..8) for a batch of positions (rank_at_cached_encoded).
void orc_rank_batch(const orc_index *ix, uint64_t n, const uint64_t *pos, uint64_t *out8) {
    for (uint64_t i = 0; i < n; ++i) {
        rankvec r = ix->ri.rank_at_cached_encoded(pos[i]);
        for (int j = 0; j < 8; ++j) out8[8 * i + j] = r[j];
    }
}

// One tag query (query_compressed / query_compressed_compact by layout). Returns number_of_runs; *n_out values written (<= cap).
uint64_t orc_tag_query(const orc_index *ix, uint64_t start, uint64_t end, uint64_t *out, uint64_t cap, uint64_t *n_out) {
    std::vector<uint64_t> v; uint64_t nr = ix->tags.query(start, end, v);
    *n_out = v.size();
    for (size_t i = 0; i < v.size() && i < cap; ++i) out[i] = v[i];
    return nr;
}

// find_mems over a batch: reads = concatenated bytes, read_off[n_reads+1]. with_tags requires loaded tags.
int orc_find_mems(const orc_index *ix, const char *reads, const uint64_t *read_off, uint64_t n_reads,
                  uint64_t min_len, uint64_t min_occ, int with_tags, int threads, int count_work, orc_result *res) {
    try {
        if (with_tags && !ix->has_tags) throw std::runtime_error("oracle: tags not loaded");
        std::memset(res, 0, sizeof(*res));
#ifdef _OPENMP
        if (threads <= 0) threads = omp_get_max_threads();
#else
        threads = 1;
#endif
        // per-chunk results keep output order = read order
        const uint64_t CH = 64;
        uint64_t n_chunks = (n_reads + CH - 1) / CH;
        struct Chunk { std::vector<MEM> mems; std::vector<uint64_t> mem_cnt, tag_cnt, tags, n_runs; };
        std::vector<Chunk> chunks(n_chunks);
        std::vector<Counters> ctrs(threads);
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
        for (int64_t ci = 0; ci < (int64_t) n_chunks; ++ci) {
#ifdef _OPENMP
            int tid = omp_get_thread_num();
#else
            int tid = 0;
#endif
            Counters *ctr = count_work ? &ctrs[tid] : nullptr;
            Chunk &ch = chunks[ci];
            std::vector<uint64_t> tv;
            uint64_t r0 = ci * CH, r1 = std::min<uint64_t>(n_reads, r0 + CH);
            for (uint64_t r = r0; r < r1; ++r) {
                size_t before = ch.mems.size();
                find_all_mems(reads + read_off[r], read_off[r + 1] - read_off[r], min_len, min_occ, ix->ri, ch.mems, ctr);
                ch.mem_cnt.push_back(ch.mems.size() - before);
                if (ctr) ctr->mems += ch.mems.size() - before;
                if (with_tags) for (size_t m = before; m < ch.mems.size(); ++m) {
                    uint64_t skipped = 0;
                    uint64_t nr = ix->tags.query(ch.mems[m].bwt_start, ch.mems[m].bwt_start + (uint64_t) ch.mems[m].size - 1, tv, &skipped);
                    ch.n_runs.push_back(nr); ch.tag_cnt.push_back(tv.size());
                    ch.tags.insert(ch.tags.end(), tv.begin(), tv.end());
                    if (ctr) { ctr->tag_runs += nr; ctr->tag_skipped += skipped; ctr->tag_item_sectors += (8 * (skipped + nr) + 31) / 32; }
                }
            }
        }
        uint64_t nm = 0, nt = 0;
        for (auto &c : chunks) { nm += c.mems.size(); nt += c.tags.size(); }
        res->n_reads = n_reads; res->n_mems = nm; res->n_tags = nt;
        res->mem_off = (uint64_t *) std::malloc((n_reads + 1) * 8);
        res->mems = (MEM *) std::malloc((nm + 1) * sizeof(MEM));
        res->tag_off = (uint64_t *) std::malloc((nm + 1) * 8);
        res->tags = (uint64_t *) std::malloc((nt + 1) * 8);
        res->n_runs = (uint64_t *) std::malloc((nm + 1) * 8);
        uint64_t mi = 0, ti = 0, ri = 0;
        for (auto &c : chunks) {
            size_t local_m = 0;
            for (size_t r = 0; r < c.mem_cnt.size(); ++r) {
                res->mem_off[ri++] = mi + local_m; local_m += c.mem_cnt[r];
            }
            for (size_t m = 0; m < c.mems.size(); ++m) {
                res->mems[mi + m] = c.mems[m];
                res->n_runs[mi + m] = with_tags ? c.n_runs[m] : 0;
            }
            size_t local_t = 0;
            for (size_t m = 0; m < c.mems.size(); ++m) {
                res->tag_off[mi + m] = ti + local_t; if (with_tags) local_t += c.tag_cnt[m];
            }
            if (!c.tags.empty()) std::memcpy(res->tags + ti, c.tags.data(), c.tags.size() * 8);
            mi += c.mems.size(); ti += c.tags.size();
        }
        res->mem_off[n_reads] = nm; res->tag_off[nm] = nt;
        Counters tot; for (auto &c : ctrs) tot.add(c);
        res->counters[0] = tot.extensions; res->counters[1] = tot.distinct_blocks; res->counters[2] = tot.mems;
        res->counters[3] = tot.tag_runs; res->counters[4] = tot.tag_skipped; res->counters[5] = tot.tag_item_sectors;
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

// query_tags over a batch (src/query_tags.cpp:87-109): per read count_encoded / count, then -- for reads with a
// match -- the tag query of the final range. One record per read: mems[i] = {0, len, range.first, range.second -
// range.first + 1} (size 0 and bwt_start 1 = the reference's empty range {1,0}); mem_off[i] = i.
int orc_query_tags(const orc_index *ix, const char *reads, const uint64_t *read_off, uint64_t n_reads,
                   int with_tags, int threads, orc_result *res) {
    try {
        if (with_tags && !ix->has_tags) throw std::runtime_error("oracle: tags not loaded");
        std::memset(res, 0, sizeof(*res));
#ifdef _OPENMP
        if (threads <= 0) threads = omp_get_max_threads();
#else
        threads = 1;
#endif
        std::vector<MEM> recs(n_reads);
        std::vector<std::vector<uint64_t>> tv(n_reads);
        std::vector<uint64_t> nruns(n_reads, 0);
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
        for (int64_t r = 0; r < (int64_t) n_reads; ++r) {
            const uint64_t len = read_off[r + 1] - read_off[r];
            auto range = ix->ri.count_encoded(reads + read_off[r], len);
            const bool hit = range.first <= range.second;
            recs[r] = MEM{0, len, hit ? range.first : 1, hit ? (int64_t) (range.second - range.first + 1) : 0};
            if (hit && with_tags) nruns[r] = ix->tags.query(range.first, range.second, tv[r]);
        }
        uint64_t nt = 0; for (auto &v : tv) nt += v.size();
        res->n_reads = n_reads; res->n_mems = n_reads; res->n_tags = nt;
        res->mem_off = (uint64_t *) std::malloc((n_reads + 1) * 8);
        res->mems = (MEM *) std::malloc((n_reads + 1) * sizeof(MEM));
        res->tag_off = (uint64_t *) std::malloc((n_reads + 1) * 8);
        res->tags = (uint64_t *) std::malloc((nt + 1) * 8);
        res->n_runs = (uint64_t *) std::malloc((n_reads + 1) * 8);
        uint64_t ti = 0;
        for (uint64_t r = 0; r < n_reads; ++r) {
            res->mem_off[r] = r; res->mems[r] = recs[r]; res->n_runs[r] = nruns[r]; res->tag_off[r] = ti;
            if (!tv[r].empty()) std::memcpy(res->tags + ti, tv[r].data(), tv[r].size() * 8);
            ti += tv[r].size();
        }
        res->mem_off[n_reads] = n_reads; res->tag_off[n_reads] = nt;
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

// stdout of query_tags for such a result: for every read with a match the two lines query_compressed_compact prints
// itself (tag_arrays.cpp:885-889) and the summary line of src/query_tags.cpp:103-108; read_index counts from first_index.
char *orc_format_query_tags(const orc_result *r, uint64_t first_index, uint64_t *len) {
    std::string out;
    for (uint64_t i = 0; i < r->n_reads; ++i) {
        const MEM &m = r->mems[i];
        if (m.size == 0) continue;
        uint64_t a = r->tag_off[i], b = r->tag_off[i + 1];
        out += "Number of unique positions: "; out += std::to_string(b - a); out += '\n';
        for (uint64_t t = a; t < b; ++t) { out += std::to_string(r->tags[t]); out += ", "; }
        out += '\n';
        out += "read_index="; out += std::to_string(first_index + i); out += "\tlen="; out += std::to_string(m.end);
        out += "\tbwt_start="; out += std::to_string(m.bwt_start); out += "\tbwt_end="; out += std::to_string(m.bwt_start + (uint64_t) m.size - 1);
        out += "\truns="; out += std::to_string(r->n_runs[i]); out += '\n';
    }
    char *buf = (char *) std::malloc(out.size() + 1);
    std::memcpy(buf, out.data(), out.size()); buf[out.size()] = 0;
    *len = out.size();
    return buf;
}

// locate_encoded for a batch of BWT ranges [first[i], second[i]]: out_off (n + 1) / out (sorted unique sequence ids per
// query, up to cap values in total); sa_out (optional, sum of range sizes values up to sa_cap): raw SA values row by row.
// Returns the total number of ids (may exceed cap: then nothing beyond cap was written).
uint64_t orc_locate_batch(const orc_index *ix, uint64_t n, const uint64_t *first, const uint64_t *second,
                          uint64_t *out_off, uint64_t *out, uint64_t cap, uint64_t *sa_out, uint64_t sa_cap) {
    uint64_t tot = 0, sa_tot = 0;
    for (uint64_t i = 0; i < n; ++i) {
        std::vector<uint64_t> sa;
        std::vector<uint64_t> v = ix->ri.locate_encoded({first[i], second[i]}, FastLocate::NO_POSITION, sa_out ? &sa : nullptr);
        out_off[i] = tot;
        for (uint64_t x : v) { if (tot < cap) out[tot] = x; ++tot; }
        for (uint64_t x : sa) { if (sa_tot < sa_cap) sa_out[sa_tot] = x; ++sa_tot; }
    }
    out_off[n] = tot;
    return tot;
}
// decompressSA, r-index.cpp:1345-1356: SA value of every BWT row through locateFirst + locateNext.
void orc_decompress_sa(const orc_index *ix, uint64_t *out) {
    uint64_t n = ix->ri.bwt_size();
    if (!n) return;
    out[0] = ix->ri.locateFirst();
    for (uint64_t i = 1; i < n; ++i) out[i] = ix->ri.locateNext(out[i - 1]);
}

void orc_result_free(orc_result *r) {
    std::free(r->mem_off); std::free(r->mems); std::free(r->tag_off); std::free(r->tags); std::free(r->n_runs);
    std::memset(r, 0, sizeof(*r));
}

// stdout text of find_mems for a result (minus the two timing lines). seq numbering starts at first_seq
// (the reference counts non-empty lines from 1). Returns a malloc'd buffer, length in *len.
char *orc_format_text(const orc_result *r, uint64_t first_seq, int with_tags, uint64_t *len) {
    std::string out;
    for (uint64_t i = 0; i < r->n_reads; ++i) {
        uint64_t a = r->mem_off[i], b = r->mem_off[i + 1];
        format_read(out, first_seq + i, r->mems + a, b - a, r->tag_off + a, r->tags, with_tags != 0);
    }
    char *buf = (char *) std::malloc(out.size() + 1);
    std::memcpy(buf, out.data(), out.size()); buf[out.size()] = 0;
    *len = out.size();
    return buf;
}
void orc_free(void *p) { std::free(p); }

int orc_max_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

} // extern "C"
