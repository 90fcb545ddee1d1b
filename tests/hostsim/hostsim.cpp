// TEST INFRASTRUCTURE ONLY — not part of the product library.
//
// Compiles the product's device arithmetic (pangenome-index_b200/csrc/pgi_core.cuh, PGI_HD empty) and
// its host flattener as plain C++ so that, on the GPU-less build box, `pytest -m "not gpu"` can check
// the flattened layout, the lane state machine and the tag plan against the oracle step by step.
// libpgi.so never contains or calls this code; the GPU tests exercise the same header through the kernels.
#include "../../pangenome-index_b200/csrc/pgi_flatten.hpp"
#include "../../pangenome-index_b200/csrc/pgi_core.cuh"
#include <cstring>
#include <algorithm>
#include <cstdlib>
#include <string>

using namespace pgi;

struct sim_index { RiImage ri; TagImage tg; bool has_tags = false; RiDev rd; TagDev td; uint32_t seed_k = 0; std::vector<SeedWord> seed; };
static thread_local std::string g_err;
// optional trace of the positions every executed extension ranks (tools/trace_blocks.py: which blocks does a batch touch?)
static std::vector<uint64_t> *g_trace = nullptr;

static void bind(sim_index *s) {
    RiDev &d = s->rd; const RiImage &im = s->ri;
    d.blocks = im.blocks.data();
    d.q_start = im.q_start.empty() ? nullptr : im.q_start.data(); d.q_val = im.q_val.empty() ? nullptr : im.q_val.data();
    d.n = im.k.n; d.n_blocks = im.k.n_blocks; d.n_quirk = im.k.n_quirk; d.bshift = im.k.bshift; d.has_quirk = im.k.has_quirk;
    for (int i = 0; i < 8; ++i) d.C[i] = im.k.C[i];
    if (s->has_tags) {
        TagDev &t = s->td; const TagImage &g = s->tg;
        t.iv.dir = g.intervals.dir.data(); t.iv.low = g.intervals.low.data();
        t.iv.size = g.intervals.k.size; t.iv.m = g.intervals.k.m; t.iv.shift = g.intervals.k.shift; t.iv.low_bytes = g.intervals.k.low_bytes;
        t.starts = g.starts.data(); t.n_starts = g.starts.size();
        t.words = g.words.data(); t.stream = g.stream.data(); t.n_items = g.n_items;
        t.width = g.width; t.layout = (uint32_t) g.layout; t.starts_regular = g.starts_regular; t.pad = 0;
    }
}

extern "C" {
const char *sim_last_error() { return g_err.c_str(); }
sim_index *sim_create(const void *ri, uint64_t ri_len, const void *tg, uint64_t tg_len, int layout) {
    try {
        sim_index *s = new sim_index();
        flatten_ri(ri, ri_len, s->ri);
        if (tg && tg_len) { flatten_tags(tg, tg_len, layout, s->tg); s->has_tags = true; }
        bind(s);
        return s;
    } catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}
void sim_free(sim_index *s) { delete s; }
// info: n, sigma, n_blocks, bucket shift, has_quirk, n_file_runs, n_ref_blocks, encoded, tag m, tag shift, tag low_bytes, starts_regular
void sim_info(const sim_index *s, uint64_t *info) {
    info[0] = s->ri.k.n; info[1] = s->ri.k.sigma; info[2] = s->ri.k.n_blocks; info[3] = s->ri.k.bshift; info[4] = s->ri.k.has_quirk;
    info[5] = s->ri.n_file_runs; info[6] = s->ri.n_ref_blocks; info[7] = s->ri.encoded;
    info[8] = s->has_tags ? s->tg.intervals.k.m : 0; info[9] = s->has_tags ? s->tg.intervals.k.shift : 0;
    info[10] = s->has_tags ? s->tg.intervals.k.low_bytes : 0; info[11] = s->has_tags ? s->tg.starts_regular : 0; info[12] = s->ri.k.n_dir_blocks; info[13] = s->ri.k.n_buckets;
}
void sim_extend_batch(const sim_index *s, uint64_t n, const uint64_t *in3, const uint8_t *sym, const uint8_t *dir, uint64_t *out3) {
    for (uint64_t i = 0; i < n; ++i) {
        BiInt b{in3[3 * i], in3[3 * i + 1], (int64_t) in3[3 * i + 2]};
        if (dir[i]) { std::swap(b.k, b.kp); b = extend_desc(s->rd, b, s->ri.k.desc_f[sym[i]]); std::swap(b.k, b.kp); }
        else b = extend_desc(s->rd, b, s->ri.k.desc_b[sym[i]]);
        out3[3 * i] = b.k; out3[3 * i + 1] = b.kp; out3[3 * i + 2] = (uint64_t) b.s;
    }
}
// Seed table (pgi_core.cuh) built by the product's own seed_entry; K = 0 removes it. Returns the K in effect (0 on an index
// with a quirk slot, where the product builds none either).
uint32_t sim_set_seed_k(sim_index *s, uint32_t K) {
    s->seed.clear(); s->seed_k = 0;
    if (!K || K > PGI_SEED_MAX_K || s->rd.has_quirk) return 0;
    s->seed.resize((size_t) 1 << (2 * K));
    for (uint32_t i = 0; i < s->seed.size(); ++i) s->seed[i] = seed_entry(s->rd, s->ri.k.desc_b, i, K);
    s->seed_k = K;
    return K;
}
// find_all_mems of one read through the lane state machine; returns the MEM count (records of 4 u64 into out, up to cap).
uint64_t sim_find_mems(const sim_index *s, const char *read, uint64_t len, uint64_t min_len, uint64_t min_occ, uint64_t *out, uint64_t cap, uint64_t *n_ext) {
    MemLane L; lane_begin(L, len, (uint32_t) min_len, s->rd.n);
    uint64_t nm = 0, ext = 0;
    // the read as the kernel sees it: bytes at offset `base` of a buffer of aligned 8-byte words (alignment varies with len)
    const uint64_t base = (len * 7 + 3) % 29;
    std::vector<uint64_t> words((base + len) / 8 + 2, 0x5A5A5A5A5A5A5A5Aull);
    std::memcpy((char *) words.data() + base, read, len);
    const uint32_t K = s->seed_k;
    // pack_reads_kernel, word for word: 2-bit codes (16 per 32-bit unit) and the "other bytes" flag of this read
    std::vector<uint32_t> codes(words.size() / 2 + 2, 0); bool other_bytes = false;
    for (size_t i = 0; i < words.size(); ++i) {
        uint32_t bad; const uint32_t c = seed_pack8(words[i], bad);
        codes[i / 2] |= c << (16 * (i & 1));
        for (uint32_t bi = 0; bi < 8; ++bi) if (((bad >> bi) & 1u) && 8 * i + bi >= base && 8 * i + bi < base + len) other_bytes = true;
    }
    uint32_t stepwise = other_bytes ? 2u : 0u;
    for (;;) {
        const int st = lane_settle(L, (uint32_t) min_len, s->rd.n);
        if (st == LANE_EMIT) {
            if (nm < cap) { out[4 * nm] = L.x; out[4 * nm + 1] = L.j; out[4 * nm + 2] = L.b.k; out[4 * nm + 3] = (uint64_t) L.b.s; }
            ++nm; lane_emitted(L, s->rd.n); continue;
        }
        if (st != LANE_EXTEND) break;
        if (K && !stepwise && lane_seed_ready(L, K, s->rd.n)) {                     // fmd_mem_kernel's seed lookup, trip for trip
            const uint64_t lo = base + L.j + 1u - K;
            const BiInt t = seed_unpack(s->seed[seed_index_packed(codes[lo >> 4], codes[(lo >> 4) + 1], lo, K)]);
            const bool ok = !ext_failed(t, min_occ);
            if (ok) { lane_seed_apply(L, t, K, (uint32_t) min_len); ext += K; }
            stepwise = ok ? 0u : 1u;
            continue;
        }
        stepwise &= 2u;
        const unsigned a = L.j < len ? (unsigned char) read[L.j] : 0;
        BiInt t = L.b;
        if (g_trace) { const uint64_t p1 = L.phase == PH_P2 ? t.kp : t.k; g_trace->push_back(p1); g_trace->push_back(p1 + (uint64_t) t.s); g_trace->push_back(L.phase); }
        if (L.phase == PH_P2) { std::swap(t.k, t.kp); t = extend_desc(s->rd, t, s->ri.k.desc_f[a]); std::swap(t.k, t.kp); }
        else t = extend_desc(s->rd, t, s->ri.k.desc_b[a]);
        ++ext;
        lane_after_ext(L, t, (uint32_t) min_len, min_occ, s->rd.n);
    }
    if (n_ext) *n_ext = ext;
    return nm;
}
// Trace control: sim_trace(1) starts recording {p1, p2, phase} per executed extension, sim_trace(0) stops; sim_trace_read copies.
void sim_trace(int on) { delete g_trace; g_trace = on ? new std::vector<uint64_t>() : nullptr; }
uint64_t sim_trace_read(uint64_t *out, uint64_t cap) {
    if (!g_trace) return 0;
    const uint64_t n = std::min<uint64_t>(cap, g_trace->size());
    if (out) std::memcpy(out, g_trace->data(), n * 8);
    return g_trace->size();
}
// Header arithmetic alone: a block with the six counts cnt[] before it and no runs, ranked at offset 0 with the descriptor
// fields tbl (pos_rank's limb dot products against Python's big integers, up to the 40-bit / 32-bit limits of the layout).
void sim_header_sums(const uint64_t *cnt, uint32_t tbl, uint64_t *WM) {
    uint32_t w[16]; for (int i = 0; i < 16; ++i) w[i] = 0;
    w[8] = 0x7777u << 16; w[9] = 0x77777777u;                           // twelve unused runs
    pgi_put_header(w, cnt);
    const PosRank r = pos_rank(w, 0u, tbl, scan_tab(tbl));
    WM[0] = r.W; WM[1] = r.M;
}
// one block as 16 u32 words (tools)
void sim_block(const sim_index *s, uint64_t b, uint32_t *w16) { for (int i = 0; i < 8; ++i) { w16[2 * i] = (uint32_t) s->ri.blocks[b * 8 + i]; w16[2 * i + 1] = (uint32_t) (s->ri.blocks[b * 8 + i] >> 32); } }
// Whole-read backward search through the count descriptors (desc_c) and extend_desc, as fmd_mem_kernel's OP 1 does:
// out2 = {range.first, size} (no match: {1, 0}). Returns 0, or 1 when the index refuses counting (count_ok == 0).
int sim_count(const sim_index *s, const char *read, uint64_t len, uint64_t *out2) {
    if (!s->ri.k.count_ok) return 1;
    BiInt b{0, 0, (int64_t) s->rd.n};
    for (uint64_t i = len; i > 0; --i) {
        const BiInt t = extend_desc(s->rd, b, s->ri.k.desc_c[(unsigned char) read[i - 1]]);
        if (ext_failed(t, 0)) { out2[0] = 1; out2[1] = 0; return 0; }
        b = t;
    }
    out2[0] = b.k; out2[1] = (uint64_t) b.s;
    return 0;
}
// SA values of rows [first, second] through locate_row / locate_next over the flattened samples (pgi_locate_batch's
// arithmetic). Returns 0, or 1 when the index refuses locate.
int sim_locate(const sim_index *s, uint64_t first, uint64_t second, uint64_t *sa_out) {
    if (!s->ri.k.locate_ok) return 1;
    LocDev l;
    auto sr = [](const SparseRankImage &im, SparseRankDev &d) { d.dir = im.dir.data(); d.low = im.low.data(); d.size = im.k.size; d.m = im.k.m; d.shift = im.k.shift; d.low_bytes = im.k.low_bytes; };
    sr(s->ri.run_starts, l.run_starts); sr(s->ri.last, l.last);
    l.samples = PackedDev{s->ri.samples.words.data(), s->ri.samples.n, s->ri.samples.width, 0};
    l.last_to_run = PackedDev{s->ri.last_to_run.words.data(), s->ri.last_to_run.n, s->ri.last_to_run.width, 0};
    l.max_length = s->ri.k.max_length;
    if (second < first) return 0;
    uint64_t v = locate_row(l, first); *sa_out++ = v;
    for (uint64_t r = first; r < second; ++r) { v = locate_next(l, v); *sa_out++ = v; }
    return 0;
}
// one tag query through tag_plan + item decoders; returns number_of_runs, *n_out sorted unique values written (<= cap)
uint64_t sim_tag_query(const sim_index *s, uint64_t start, uint64_t end, uint64_t *out, uint64_t cap, uint64_t *n_out) {
    TagPlan p = tag_plan(s->td, start, end);
    std::vector<uint64_t> v; v.reserve(p.count);
    if (s->td.layout == 0) for (uint64_t i = 0; i < p.count; ++i) v.push_back(tag_item_compact(s->td, p.first_item + p.skip + i));
    else {
        uint64_t loc = p.first_item;
        for (uint64_t i = 0; i < p.skip; ++i) (void) tag_bytecode_next(s->td, loc);
        for (uint64_t i = 0; i < p.count; ++i) v.push_back(tag_value_bytecode(tag_bytecode_next(s->td, loc)));
    }
    std::sort(v.begin(), v.end()); v.erase(std::unique(v.begin(), v.end()), v.end());
    *n_out = v.size();
    for (size_t i = 0; i < v.size() && i < cap; ++i) out[i] = v[i];
    return p.n_runs;
}
}
