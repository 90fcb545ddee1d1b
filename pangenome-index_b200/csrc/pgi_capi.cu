// pgi_capi.cu — the C ABI of include/pgi.h: index upload, batch contexts and the launch pipeline.
// Everything a caller can reach goes through here; there is no CPU implementation of any query.
#include <cub/device/device_scan.cuh>
#include <cub/device/device_segmented_sort.cuh>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <sched.h>
#include <thread>
#include <string>
#include <vector>
#include "../../include/pgi.h"
#include "pgi_flatten.hpp"
#include "pgi_image.hpp"
#include "pgi_tag_kernels.cuh"
#include "pgi_build.cuh"

using namespace pgi;

namespace {

thread_local std::string g_err;
int fail(int code, const std::string &msg) { g_err = msg; return code; }

struct CudaError { cudaError_t e; const char *what; };
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) throw CudaError{e_, #call}; } while (0)

// Grow-only device / pinned-host buffers.
struct DevBuf {
    void *p = nullptr; size_t cap = 0;
    void need(size_t bytes) {
        if (bytes <= cap) return;
        if (p) CK(cudaFree(p));
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        CK(cudaMalloc(&p, want)); cap = want;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <typename T> T *as() const { return (T *) p; }
};
struct PinBuf {
    void *p = nullptr; size_t cap = 0;
    void need(size_t bytes) {
        if (bytes <= cap) return;
        if (p) CK(cudaFreeHost(p));
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        CK(cudaHostAlloc(&p, want, cudaHostAllocDefault)); cap = want;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
    template <typename T> T *as() const { return (T *) p; }
};

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); CK(cudaSetDevice(dev)); }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

template <typename T> T *upload(const std::vector<T> &v, std::vector<void *> &owned, uint64_t &bytes, size_t pad_elems = 0) {
    size_t n = v.size() + pad_elems;
    if (n == 0) n = 1;
    void *d = nullptr;
    CK(cudaMalloc(&d, n * sizeof(T)));
    owned.push_back(d); bytes += n * sizeof(T);
    CK(cudaMemset(d, 0, n * sizeof(T)));
    if (!v.empty()) CK(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return (T *) d;
}

} // namespace

struct pgi_index {
    int device = 0;
    RiDev ri; TagDev tg; LocDev loc; bool has_tags = false, has_ri = true, has_locate = false;
    uint64_t n_rs_dir = 0, n_rs_low = 0, n_last_dir = 0, n_last_low = 0, n_samples_words = 0, n_ltr_words = 0;
    RiConsts consts; SparseRankConsts ivk;
    uint4 *d_seed = nullptr; uint32_t seed_k = 0;          // seed table (pgi_core.cuh): 4^seed_k entries, built on the device at load
    uint32_t *d_desc = nullptr, *d_desc_c = nullptr;      // extension descriptors; LF-step descriptors of the whole-read search
    std::vector<void *> owned;
    std::vector<void *> owned_host;                        // sections kept in pinned host memory (out-of-core mode)
    uint64_t host_resident_bytes = 0;
    // sizes of the uploaded arrays, for clone
    uint64_t n_block_words = 0, n_quirk = 0, n_iv_dir = 0, n_iv_low = 0, n_starts = 0, n_words = 0, n_stream = 0;
    pgi_index_info_t info;
    std::mutex compute_mu;                                  // see pgi_find_mems
    // How the steps that concurrent callers / in-flight batches enqueue on this index share the GPU (batch_enqueue):
    // ev_end = end of the step enqueued last, ev_k1 = end of its fmd_mem_kernel launch; written under compute_mu.
    cudaEvent_t ev_end = nullptr, ev_k1 = nullptr; bool ev_end_set = false, ev_k1_set = false;
    std::mutex upload_mu, download_mu;                      // one bulk copy per direction at a time (pgi_find_mems)
    std::mutex pool_mu; std::vector<pgi_batch *> pool;     // idle batch contexts used by pgi_find_mems
    int sm_count = 0;
};

extern "C" const char *pgi_last_error(void) { return g_err.c_str(); }

extern "C" int pgi_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

static int check_device(int device) {
    int n = pgi_device_count();
    if (n <= 0) return fail(PGI_ENODEVICE, "no usable CUDA device: libpgi has no CPU path");
    if (device < 0 || device >= n) return fail(PGI_EINVAL, "device index out of range");
    return 0;
}

static void fill_info(pgi_index *ix, const ImageMeta &m, uint64_t bytes) {
    pgi_index_info_t &f = ix->info; std::memset(&f, 0, sizeof(f));
    f.n = m.k.n; f.sigma = m.k.sigma; f.is_encoded = m.encoded; f.n_ref_blocks = m.n_ref_blocks; f.n_runs = m.n_file_runs;
    f.n_dev_blocks = m.k.n_blocks; f.bucket_shift = m.k.bshift; f.n_crowded_buckets = m.k.n_dir_blocks; f.has_quirk_slot = m.k.has_quirk;
    if (m.has_tags) {
        f.has_tags = 1; f.tag_layout = m.tag_layout; f.tag_runs = m.iv_k.m; f.tag_items = m.tag_n_items;
        f.tag_item_width = m.tag_width; f.tag_universe = m.iv_k.size;
    }
    f.device_bytes = bytes; f.host_resident_bytes = ix->host_resident_bytes; f.has_locate = m.k.locate_ok; f.max_length = m.k.max_length; f.seed_k = ix->seed_k;
    for (int i = 0; i < 8; ++i) f.C[i] = m.k.C[i];
    std::memcpy(f.sym_map, m.k.sym_map, 256);
}

// Host -> HBM copy of one section. chunk == 0: one cudaMemcpy (the source is ordinary process memory). chunk > 0: the
// source is a file mapping; it is copied through two pinned staging buffers of `chunk` bytes, the memcpy out of the page
// cache (page faults, read-ahead) of chunk i + 1 overlapping the DMA of chunk i. Host memory stays at 2 x chunk however
// large the index is -- the role of chunk_size_mb in the reference's ChunkedMemoryMappedIndexManager
// (src/find_mems_chunked.cpp:29-43).
struct Stager {
    size_t chunk = 0; void *pin[2] = {nullptr, nullptr}; cudaEvent_t done[2]; cudaStream_t st = nullptr; int turn = 0; bool ok = false;
    uint64_t chunks = 0;
    void init(size_t chunk_bytes) {
        chunk = chunk_bytes;
        if (!chunk) return;
        CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) { CK(cudaHostAlloc(&pin[i], chunk, cudaHostAllocDefault)); CK(cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming)); }
        ok = true;
    }
    void copy(void *dst, const void *src, size_t bytes) {
        if (!chunk) { CK(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice)); return; }
        for (size_t at = 0; at < bytes; at += chunk) {
            const size_t n = std::min(chunk, bytes - at);
            CK(cudaEventSynchronize(done[turn]));                       // the buffer's previous DMA has drained
            std::memcpy(pin[turn], (const char *) src + at, n);
            CK(cudaMemcpyAsync((char *) dst + at, pin[turn], n, cudaMemcpyHostToDevice, st));
            CK(cudaEventRecord(done[turn], st));
            turn ^= 1; ++chunks;
        }
    }
    void finish() { if (chunk && st) CK(cudaStreamSynchronize(st)); }
    ~Stager() {
        if (!ok) { if (st) cudaStreamDestroy(st); for (int i = 0; i < 2; ++i) if (pin[i]) cudaFreeHost(pin[i]); return; }
        cudaStreamSynchronize(st);
        for (int i = 0; i < 2; ++i) { cudaFreeHost(pin[i]); cudaEventDestroy(done[i]); }
        cudaStreamDestroy(st);
    }
};

// Seed table (pgi_core.cuh): one thread per K-mer runs the K backward extensions from the full interval.
__global__ void seed_table_kernel(RiDev ix, const uint32_t *desc, uint32_t K, uint32_t n_entries, uint4 *out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_entries) return;
    const SeedWord e = seed_entry(ix, desc, i, K);
    out[i] = make_uint4(e.x, e.y, e.z, e.w);
}
// K by index size: a K-mer should still occur (4^K well below n) and the table should stay a small fraction of the index.
// PGI_SEEDK=<0..12> overrides it (0: no table) for tuning runs and tests. No table for indexes with a quirk slot
// (interval sizes are not monotone there) or without an r-index part.
static uint32_t choose_seed_k(const ImageMeta &m) {
    if (!m.has_ri || m.k.has_quirk || m.k.n < 16) return 0;
    // (config 4, n = 2^36.2: K = 12 / 13 / 14 / 15 -> 10.64 / 10.17 / 10.92 / 12.33 ms per step; config 3, n = 2^32.6: 12 beats 14)
    uint32_t K = m.k.n >= (1ull << 35) ? 13u : m.k.n >= (1ull << 28) ? 12u : m.k.n >= (1ull << 22) ? 10u : m.k.n >= (1ull << 16) ? 8u : 6u;
    if (const char *e = std::getenv("PGI_SEEDK")) { const long v = std::strtol(e, nullptr, 10); if (v >= 0 && v <= (long) PGI_SEED_MAX_K) K = (uint32_t) v; }
    return K;
}
static void build_seed_table(pgi_index *ix, uint32_t K, uint64_t &bytes) {
    ix->seed_k = 0; ix->d_seed = nullptr;
    if (!K) return;
    const uint32_t n_entries = 1u << (2u * K);
    void *d = nullptr; CK(cudaMalloc(&d, (size_t) n_entries * 16)); ix->owned.push_back(d); bytes += (uint64_t) n_entries * 16;
    seed_table_kernel<<<(n_entries + 255) / 256, 256>>>(ix->ri, ix->d_desc, K, n_entries, (uint4 *) d);
    CK(cudaGetLastError());
    ix->d_seed = (uint4 *) d; ix->seed_k = K;
}

// Allocate + fill the device copy of every section of a flattened index (fresh from flatten_ri / flatten_tags, or a
// mapped image file) and wire the device structs. Throws CudaError.
static pgi_index *upload_view(const IndexView &v, int device, size_t chunk_bytes, uint64_t *chunks_out) {
    DeviceGuard g(device);
    pgi_index *ix = new pgi_index();
    try {
        const ImageMeta &m = v.m;
        ix->device = device; ix->consts = m.k; ix->has_ri = m.has_ri != 0;
        cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, device)); ix->sm_count = prop.multiProcessorCount;
        uint64_t bytes = 0;
        Stager stg; stg.init(chunk_bytes);
        // Out-of-core mode (second half; the role of the reference's ChunkedMemoryMappedIndexManager for an
        // index that does not fit the device): the run-block array -- by far the largest section -- stays in PINNED HOST memory,
        // mapped into the device's address space, and the kernels read it over PCIe (zero copy) with the same code. Chosen when
        // the section does not fit the free HBM (leaving 1/8 of it for work buffers) or when PGI_HOST_BLOCKS=1 forces it (tests).
        // Functional, not fast: every block fetch becomes a 64-byte PCIe read.
        size_t free_b = 0, total_b = 0; CK(cudaMemGetInfo(&free_b, &total_b));
        const char *force = std::getenv("PGI_HOST_BLOCKS");
        auto up = [&](int s, uint64_t &n_elems) -> void * {
            const uint64_t nb = v.bytes[s]; n_elems = nb / kSectionElem[s];
            void *d = nullptr; const size_t alloc = nb ? (size_t) nb : kSectionElem[s];
            const bool on_host = s == SEC_BLOCKS && nb && ((force && force[0] == '1') || nb > free_b - free_b / 8);
            if (on_host) {
                CK(cudaHostAlloc(&d, alloc, cudaHostAllocMapped | cudaHostAllocPortable)); ix->owned_host.push_back(d); ix->host_resident_bytes += alloc;
                std::memcpy(d, v.ptr[s], (size_t) nb);
                void *dp = nullptr; CK(cudaHostGetDevicePointer(&dp, d, 0));
                return dp;
            }
            CK(cudaMalloc(&d, alloc)); ix->owned.push_back(d); bytes += alloc;
            if (nb) stg.copy(d, v.ptr[s], (size_t) nb); else CK(cudaMemset(d, 0, alloc));
            return d;
        };
        RiDev &d = ix->ri; uint64_t nq2 = 0;
        d.blocks = (const uint64_t *) up(SEC_BLOCKS, ix->n_block_words);
        d.q_start = (const uint64_t *) up(SEC_Q_START, ix->n_quirk); d.q_val = (const uint64_t *) up(SEC_Q_VAL, nq2);
        d.n = m.k.n; d.n_blocks = m.k.n_blocks; d.n_quirk = m.k.n_quirk; d.bshift = m.k.bshift; d.has_quirk = m.k.has_quirk; d.n_buckets_dbg = m.k.n_buckets;
        for (int i = 0; i < 8; ++i) d.C[i] = m.k.C[i];
        std::vector<uint32_t> desc(PGI_DESC_WORDS);      // 512 descriptors, then their 512 x 4 scan tables (pos_rank)
        for (int i = 0; i < 256; ++i) { desc[i] = m.k.desc_b[i]; desc[256 + i] = m.k.desc_f[i]; }
        for (int i = 0; i < 512; ++i) {
            const ScanTab t = scan_tab(desc[i] & 0xFFFFFFu);
            desc[512 + 4 * i] = t.x; desc[512 + 4 * i + 1] = t.y; desc[512 + 4 * i + 2] = t.z; desc[512 + 4 * i + 3] = t.w;
        }
        ix->d_desc = upload(desc, ix->owned, bytes);
        for (int i = 0; i < 512; ++i) {                    // the count bank in the same shape (both halves: the kernel never asks for "forward")
            desc[i] = m.k.desc_c[i & 255];
            const ScanTab t = scan_tab(desc[i] & 0xFFFFFFu);
            desc[512 + 4 * i] = t.x; desc[512 + 4 * i + 1] = t.y; desc[512 + 4 * i + 2] = t.z; desc[512 + 4 * i + 3] = t.w;
        }
        ix->d_desc_c = upload(desc, ix->owned, bytes);
        std::memset(&ix->loc, 0, sizeof(ix->loc));
        if (m.has_ri && m.has_locate) {
            LocDev &l = ix->loc; ix->has_locate = true;
            auto up_sr = [&](const SparseRankConsts &k, int sdir, int slow, SparseRankDev &sd, uint64_t &nd, uint64_t &nl) {
                sd.dir = (const uint64_t *) up(sdir, nd); sd.low = (const uint8_t *) up(slow, nl);
                sd.size = k.size; sd.m = k.m; sd.shift = k.shift; sd.low_bytes = k.low_bytes;
            };
            up_sr(m.rs_k, SEC_RS_DIR, SEC_RS_LOW, l.run_starts, ix->n_rs_dir, ix->n_rs_low);
            up_sr(m.last_k, SEC_LAST_DIR, SEC_LAST_LOW, l.last, ix->n_last_dir, ix->n_last_low);
            l.samples.words = (const uint64_t *) up(SEC_SAMPLES, ix->n_samples_words);
            l.samples.n = m.samples_n; l.samples.width = (uint32_t) m.samples_width; l.samples.pad = 0;
            l.last_to_run.words = (const uint64_t *) up(SEC_LAST_TO_RUN, ix->n_ltr_words);
            l.last_to_run.n = m.ltr_n; l.last_to_run.width = (uint32_t) m.ltr_width; l.last_to_run.pad = 0;
            l.max_length = m.k.max_length;
        }
        std::memset(&ix->tg, 0, sizeof(ix->tg));
        if (m.has_tags) {
            TagDev &t = ix->tg; ix->has_tags = true; ix->ivk = m.iv_k;
            t.iv.dir = (const uint64_t *) up(SEC_IV_DIR, ix->n_iv_dir); t.iv.low = (const uint8_t *) up(SEC_IV_LOW, ix->n_iv_low);
            t.iv.size = m.iv_k.size; t.iv.m = m.iv_k.m; t.iv.shift = m.iv_k.shift; t.iv.low_bytes = m.iv_k.low_bytes;
            t.starts = (const uint64_t *) up(SEC_STARTS, ix->n_starts); t.n_starts = ix->n_starts;
            t.words = (const uint64_t *) up(SEC_WORDS, ix->n_words);
            t.stream = (const uint8_t *) up(SEC_STREAM, ix->n_stream);
            t.n_items = m.tag_n_items; t.width = (uint32_t) m.tag_width; t.layout = (uint32_t) m.tag_layout; t.starts_regular = (uint32_t) m.tag_starts_regular; t.pad = 0;
        }
        stg.finish();
        build_seed_table(ix, choose_seed_k(m), bytes);
        fill_info(ix, m, bytes);
        CK(cudaDeviceSynchronize());
        if (chunks_out) *chunks_out = stg.chunks;
        return ix;
    } catch (...) { pgi_index_free(ix); throw; }
}

// Flatten the SDSL files on the host (no device involved).
static void flatten_files(const void *ri_bytes, size_t ri_len, const void *tag_bytes, size_t tag_len, int tag_layout,
                          RiImage &ri, TagImage &tg, IndexView &v) {
    const bool has_tags = tag_bytes && tag_len;
    std::memset(&ri.k, 0, sizeof(ri.k));
    if (ri_bytes) flatten_ri(ri_bytes, ri_len, ri);
    if (has_tags) flatten_tags(tag_bytes, tag_len, tag_layout, tg);
    view_of(ri_bytes ? &ri : nullptr, has_tags ? &tg : nullptr, v);
}

#define PGI_CATCH_ALL \
    catch (const FormatError &e) { return fail(PGI_EFORMAT, e.what()); } \
    catch (const CudaError &e) { return fail(PGI_ECUDA, std::string(e.what) + ": " + cudaGetErrorString(e.e)); } \
    catch (const std::bad_alloc &) { return fail(PGI_ENOMEM, "out of host memory while flattening the index"); } \
    catch (const std::exception &e) { return fail(PGI_EINVAL, e.what()); }

extern "C" int pgi_index_create(const void *ri_bytes, size_t ri_len, const void *tag_bytes, size_t tag_len,
                                int tag_layout, int device, pgi_index **out) {
    if ((!ri_bytes && !tag_bytes) || !out) return fail(PGI_EINVAL, "null argument");
    if (int rc = check_device(device)) return rc;
    try {
        RiImage ri; TagImage tg; IndexView v;
        flatten_files(ri_bytes, ri_len, tag_bytes, tag_len, tag_layout, ri, tg, v);
        *out = upload_view(v, device, 0, nullptr);
        return PGI_OK;
    } PGI_CATCH_ALL
}

// ------------------------------------------------------------------------------------------ GPU-side construction
namespace {
template <typename T> T *dev_alloc(uint64_t n, std::vector<void *> &owned, uint64_t &bytes, bool zero = false) {
    void *d = nullptr; const size_t b = (size_t) std::max<uint64_t>(1, n) * sizeof(T);
    CK(cudaMalloc(&d, b)); owned.push_back(d); bytes += b;
    if (zero) CK(cudaMemset(d, 0, b));
    return (T *) d;
}
struct Tmp {                        // scratch device buffers of one construction, freed on scope exit
    std::vector<void *> v;
    template <typename T> T *get(uint64_t n) { void *d = nullptr; CK(cudaMalloc(&d, (size_t) std::max<uint64_t>(1, n) * sizeof(T))); v.push_back(d); return (T *) d; }
    void drop(const void *p) { for (auto &x : v) if (x == p) { cudaFree(x); x = nullptr; } }
    ~Tmp() { for (void *x : v) if (x) cudaFree(x); }
};
inline unsigned grid_of(uint64_t n, unsigned threads) { return (unsigned) std::max<uint64_t>(1, (n + threads - 1) / threads); }

// The descriptor banks + seed table + info of an index whose device arrays are in place (shared by both constructions).
void finish_index(pgi_index *ix, const ImageMeta &m, uint64_t &bytes) {
    std::vector<uint32_t> desc(PGI_DESC_WORDS);
    auto bank = [&](const uint32_t *b, const uint32_t *f) {
        for (int i = 0; i < 256; ++i) { desc[i] = b[i]; desc[256 + i] = f[i]; }
        for (int i = 0; i < 512; ++i) { const ScanTab t = scan_tab(desc[i] & 0xFFFFFFu); desc[512 + 4 * i] = t.x; desc[512 + 4 * i + 1] = t.y; desc[512 + 4 * i + 2] = t.z; desc[512 + 4 * i + 3] = t.w; }
        return upload(desc, ix->owned, bytes);
    };
    ix->d_desc = bank(m.k.desc_b, m.k.desc_f);
    ix->d_desc_c = bank(m.k.desc_c, m.k.desc_c);
    build_seed_table(ix, choose_seed_k(m), bytes);
    fill_info(ix, m, bytes);
}
} // namespace

extern "C" int pgi_index_create_from_runs(const uint8_t *sym, const uint64_t *len, uint64_t n_runs,
                                          const uint64_t *tag_val, const uint64_t *tag_len, uint64_t n_tag_runs,
                                          int device_pointers, int device, pgi_index **out) {
    if (!sym || !len || !n_runs || !out || (n_tag_runs && (!tag_val || !tag_len))) return fail(PGI_EINVAL, "null argument");
    if (int rc = check_device(device)) return rc;
    if (n_runs >> 32) return fail(PGI_EINVAL, "more than 2^32 BWT runs");
    pgi_index *ix = nullptr;
    try {
        DeviceGuard g(device);
        namespace B = pgi::build;
        auto pol = thrust::device;
        Tmp tmp;
        ix = new pgi_index(); ix->device = device;
        cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, device)); ix->sm_count = prop.multiProcessorCount;
        uint64_t bytes = 0;
        ImageMeta m; std::memset(&m, 0, sizeof(m));
        // ---- inputs on the device
        const uint8_t *d_sym = sym; const uint64_t *d_len = len, *d_tv = tag_val, *d_tl = tag_len;
        if (!device_pointers) {
            uint8_t *a = tmp.get<uint8_t>(n_runs); uint64_t *b = tmp.get<uint64_t>(n_runs);
            CK(cudaMemcpy(a, sym, n_runs, cudaMemcpyHostToDevice)); CK(cudaMemcpy(b, len, n_runs * 8, cudaMemcpyHostToDevice)); d_sym = a; d_len = b;
            if (n_tag_runs) {
                uint64_t *c = tmp.get<uint64_t>(n_tag_runs), *d = tmp.get<uint64_t>(n_tag_runs);
                CK(cudaMemcpy(c, tag_val, n_tag_runs * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d, tag_len, n_tag_runs * 8, cudaMemcpyHostToDevice)); d_tv = c; d_tl = d;
            }
        }
        thrust::counting_iterator<uint64_t> idx(0);
        // ---- symbol codes, validation, totals, logical runs (an end marker record of length f is f runs, r-index.cpp:840-848)
        uint8_t *code0 = tmp.get<uint8_t>(n_runs);
        thrust::transform(pol, d_sym, d_sym + n_runs, code0, B::CodeOf());
        if (thrust::transform_reduce(pol, idx, idx + n_runs, B::BadRun{code0, d_len}, 0ull, thrust::plus<uint64_t>()))
            throw FormatError("BWT run with a symbol outside {\\n,A,C,G,N,T} or of length 0");
        uint64_t totals[6];
        for (uint8_t c = 0; c < 6; ++c) totals[c] = thrust::transform_reduce(pol, idx, idx + n_runs, B::LenIfCode{c, code0, d_len}, 0ull, thrust::plus<uint64_t>());
        const uint64_t logical_runs = thrust::transform_reduce(pol, idx, idx + n_runs, B::LogicalRuns{code0, d_len}, 0ull, thrust::plus<uint64_t>());
        uint64_t n = 0; for (uint64_t t : totals) n += t;
        if (totals[1] >> 40 || totals[2] >> 40 || totals[3] >> 40 || totals[5] >> 40) throw FormatError("index longer than 2^40 positions is not supported by the device layout");
        if (totals[0] >> 32 || totals[4] >> 32) throw FormatError("more than 2^32 endmarkers or N symbols is not supported by the device layout");
        // ---- maximal runs and their start positions
        uint8_t *code = tmp.get<uint8_t>(n_runs); uint64_t *rlen = tmp.get<uint64_t>(n_runs);
        const uint64_t r = (uint64_t) (thrust::reduce_by_key(pol, code0, code0 + n_runs, d_len, code, rlen).first - code);
        tmp.drop(code0);
        uint64_t *starts = tmp.get<uint64_t>(r + 1);
        thrust::exclusive_scan(pol, rlen, rlen + r, starts);
        CK(cudaMemcpy(starts + r, &n, 8, cudaMemcpyHostToDevice));
        // ---- bucket shift: the largest s <= 8 with at most 1/64 of the buckets crowded (pgi_flatten.hpp build_buckets)
        unsigned S = PGI_MAX_BSHIFT; bool forced = false;
        if (const char *e = std::getenv("PGI_BSHIFT")) {
            S = (unsigned) std::atoi(e); forced = true;
            if (S < PGI_MIN_BSHIFT || S > PGI_MAX_BSHIFT) throw FormatError("PGI_BSHIFT outside [4, 12]");
        }
        unsigned long long *d_ctr = tmp.get<unsigned long long>(1);
        uint32_t *frun = nullptr; uint64_t nb = 0; unsigned long long n_crowded = 0;
        for (;; --S) {
            nb = (n >> S) + 1;
            if (frun) tmp.drop(frun);
            frun = tmp.get<uint32_t>(nb + 1);
            // frun[b] = (number of run starts <= b << S) - 1: the run holding the first position of bucket b
            auto q = thrust::make_transform_iterator(idx, B::ShiftLeft{S});
            thrust::upper_bound(pol, starts, starts + r, q, q + nb, frun);
            thrust::transform(pol, frun, frun + nb, frun, B::MinusOne());
            CK(cudaMemset(d_ctr, 0, 8));
            B::count_entries_kernel<<<grid_of(nb, 256), 256>>>(frun, starts, rlen, nb, S, r, n, nullptr, nullptr, d_ctr);
            CK(cudaGetLastError());
            CK(cudaMemcpy(&n_crowded, d_ctr, 8, cudaMemcpyDeviceToHost));
            if (forced || S == PGI_MIN_BSHIFT || n_crowded <= ((n >> S) + 1) / crowded_div()) break;
        }
        uint32_t *n_chunks = tmp.get<uint32_t>(nb + 1); uint64_t *chunk_base = tmp.get<uint64_t>(nb + 1); uint8_t *sub_shift = tmp.get<uint8_t>(nb + 1);
        CK(cudaMemset(n_chunks, 0, (nb + 1) * 4)); CK(cudaMemset(d_ctr, 0, 8));
        B::count_entries_kernel<<<grid_of(nb, 256), 256>>>(frun, starts, rlen, nb, S, r, n, n_chunks, sub_shift, d_ctr);
        CK(cudaGetLastError());
        thrust::exclusive_scan(pol, thrust::make_transform_iterator(n_chunks, B::Widen()), thrust::make_transform_iterator(n_chunks, B::Widen()) + nb + 1, chunk_base);
        uint64_t total_chunks = 0; CK(cudaMemcpy(&total_chunks, chunk_base + nb, 8, cudaMemcpyDeviceToHost));
        const uint64_t n_blocks = nb + total_chunks;
        if (n_blocks >= 0xFFFFFFFFull) throw FormatError("too many run blocks (>= 2^32 - 1) for the 32-bit block indices of the MEM kernel");
        uint32_t *blocks = dev_alloc<uint32_t>(n_blocks * PGI_BLOCK_U32, ix->owned, bytes);
        B::bucket_kernel<<<grid_of(nb, 128), 128>>>(frun, starts, rlen, code, nb, S, r, n, n_chunks, sub_shift, chunk_base, blocks);
        CK(cudaGetLastError());
        {   // cumulative headers: field by field, the in-bucket totals sit in the primary blocks themselves
            uint64_t *before = tmp.get<uint64_t>(nb);
            for (unsigned field = 1; field <= 6; ++field) {
                auto in = thrust::make_transform_iterator(idx, B::FieldOfBlock{blocks, field});
                thrust::exclusive_scan(pol, in, in + nb, before);
                B::header_kernel<<<grid_of(nb, 256), 256>>>(blocks, before, nb, field, n_chunks, chunk_base);
                CK(cudaGetLastError());
            }
            B::pack_headers_kernel<<<grid_of(n_blocks, 256), 256>>>(blocks, n_blocks);
            CK(cudaGetLastError());
            CK(cudaDeviceSynchronize());
            tmp.drop(before);
        }
        tmp.drop(frun); tmp.drop(n_chunks); tmp.drop(chunk_base); tmp.drop(sub_shift); tmp.drop(starts); tmp.drop(code); tmp.drop(rlen);
        // ---- scalars and tables, as FastLocate::calculate_C (r-index.hpp:440-482) + serialize_encoded / load_encoded leave them
        RiImage im; RiConsts &k = im.k; std::memset(&k, 0, sizeof(k));
        bool present[6]; unsigned sigma = 0;
        for (unsigned c = 0; c < 6; ++c) { present[c] = totals[c] != 0; if (present[c]) { k.sym_map[kNuc[c]] = (uint8_t) sigma; k.C[sigma] = 0; ++sigma; } }
        { uint64_t acc = 0; unsigned j = 0; for (unsigned c = 0; c < 6; ++c) if (present[c]) { k.C[j++] = acc; acc += totals[c]; } }
        if (!present[0]) throw FormatError("a BWT without end markers is not a multi-string BWT");
        if (sigma != (present[4] ? 6u : 5u)) throw FormatError("encoded r-index whose alphabet size disagrees with its N flag (unreadable by the reference too)");
        k.n = n; k.sigma = sigma; k.n_blocks = n_blocks; k.bshift = S; k.n_buckets = nb; k.n_dir_blocks = n_crowded; k.n_quirk = 0; k.mem_tail = 0; k.max_length = 1;
        build_descriptors(im, present, false);
        k.locate_ok = 0;
        m.k = k; m.n_file_runs = logical_runs; m.n_ref_blocks = logical_runs / 10 + (logical_runs % 10 != 0); m.encoded = 1; m.has_ri = 1;
        ix->consts = k; ix->has_ri = true; ix->n_block_words = n_blocks * PGI_BLOCK_WORDS;
        RiDev &d = ix->ri; std::memset(&d, 0, sizeof(d));
        d.blocks = (const uint64_t *) blocks; d.q_start = dev_alloc<uint64_t>(1, ix->owned, bytes, true); d.q_val = dev_alloc<uint64_t>(1, ix->owned, bytes, true);
        d.n = n; d.n_blocks = n_blocks; d.n_quirk = 0; d.bshift = S; d.has_quirk = 0; d.n_buckets_dbg = nb;
        for (int i = 0; i < 8; ++i) d.C[i] = k.C[i];
        std::memset(&ix->loc, 0, sizeof(ix->loc)); std::memset(&ix->tg, 0, sizeof(ix->tg));
        // ---- tag arrays (sdsl-compact semantics: prep_tags_from_runs / merge_tags' writer + load_compressed_tags_compact)
        if (n_tag_runs) {
            if (thrust::transform_reduce(pol, idx, idx + n_tag_runs, B::ZeroLen{d_tl}, 0ull, thrust::plus<uint64_t>())) throw FormatError("tag run of length 0");
            uint64_t *run_start = tmp.get<uint64_t>(n_tag_runs + 1), *item_off = tmp.get<uint64_t>(n_tag_runs + 1);
            thrust::exclusive_scan(pol, d_tl, d_tl + n_tag_runs, run_start);
            auto pc = thrust::make_transform_iterator(idx, B::Pieces{d_tl});
            thrust::exclusive_scan(pol, pc, pc + n_tag_runs, item_off);
            const uint64_t total_len = thrust::reduce(pol, d_tl, d_tl + n_tag_runs, 0ull, thrust::plus<uint64_t>());
            const uint64_t n_items = thrust::reduce(pol, pc, pc + n_tag_runs, 0ull, thrust::plus<uint64_t>());
            auto nd = thrust::make_transform_iterator(idx, B::NodeOf{d_tv});
            const uint64_t max_node = thrust::reduce(pol, nd, nd + n_tag_runs, 0ull, thrust::maximum<uint64_t>());
            unsigned hi = 0; while (hi < 63 && (max_node >> (hi + 1))) ++hi;
            const unsigned width = 11 + hi + 1;
            if (width > 64) throw FormatError("tag item width outside [1,64]");
            uint64_t *item_val = tmp.get<uint64_t>(n_items), *item_pos = tmp.get<uint64_t>(n_items);
            B::tag_items_kernel<<<grid_of(n_tag_runs, 256), 256>>>(d_tv, d_tl, run_start, item_off, n_tag_runs, item_val, item_pos);
            CK(cudaGetLastError());
            tmp.drop(run_start); tmp.drop(item_off);
            const uint64_t nw = (n_items * width + 63) / 64;
            uint64_t *words = dev_alloc<uint64_t>(nw + 2, ix->owned, bytes, true);
            B::pack_items_kernel<<<grid_of(nw, 256), 256>>>(item_val, n_items, width, nw, words);
            CK(cudaGetLastError()); CK(cudaDeviceSynchronize());
            tmp.drop(item_val);
            // bwt_intervals: universe total_len + 1, one 1 at the first BWT position of every item (build_sparse_rank_from)
            SparseRankConsts &sk = m.iv_k; std::memset(&sk, 0, sizeof(sk));
            sk.size = total_len + 1; sk.m = n_items;
            unsigned sh = 0; while (sh < 62 && (sk.size >> sh) > n_items / 4 + 1) ++sh;
            sk.shift = sh; sk.low_bytes = sh <= 8 ? 1 : sh <= 16 ? 2 : sh <= 32 ? 4 : 8; sk.n_buckets = (sk.size >> sh) + 1;
            uint64_t *dir = dev_alloc<uint64_t>(sk.n_buckets + 1, ix->owned, bytes);
            auto qb = thrust::make_transform_iterator(idx, B::ShiftLeft{sh});
            thrust::lower_bound(pol, item_pos, item_pos + n_items, qb, qb + sk.n_buckets + 1, dir);       // dir[b] = items starting before b << shift
            uint8_t *low = dev_alloc<uint8_t>(n_items * sk.low_bytes + 16, ix->owned, bytes, true);
            B::low_parts_kernel<<<grid_of(n_items, 256), 256>>>(item_pos, n_items, sh, sk.low_bytes, low);
            CK(cudaGetLastError());
            const uint64_t n_starts = (n_items + 9) / 10;
            uint64_t *st = dev_alloc<uint64_t>(n_starts, ix->owned, bytes);
            B::iota10_kernel<<<grid_of(n_starts, 256), 256>>>(n_starts, st);
            CK(cudaGetLastError()); CK(cudaDeviceSynchronize());
            tmp.drop(item_pos);
            TagDev &t = ix->tg; ix->has_tags = true; ix->ivk = sk;
            t.iv.dir = dir; t.iv.low = low; t.iv.size = sk.size; t.iv.m = sk.m; t.iv.shift = sk.shift; t.iv.low_bytes = sk.low_bytes;
            t.starts = st; t.n_starts = n_starts; t.words = words; t.stream = dev_alloc<uint8_t>(1, ix->owned, bytes, true);
            t.n_items = n_items; t.width = width; t.layout = 0; t.starts_regular = 1; t.pad = 0;
            ix->n_iv_dir = sk.n_buckets + 1; ix->n_iv_low = n_items * sk.low_bytes + 16; ix->n_starts = n_starts; ix->n_words = nw + 2; ix->n_stream = 1;
            m.has_tags = 1; m.tag_layout = 0; m.tag_n_items = n_items; m.tag_width = width; m.tag_starts_regular = 1;
        }
        ix->n_quirk = 1;
        finish_index(ix, m, bytes);
        CK(cudaDeviceSynchronize());
        *out = ix; return PGI_OK;
    } catch (const FormatError &e) { pgi_index_free(ix); return fail(PGI_EFORMAT, e.what()); }
      catch (const CudaError &e) { pgi_index_free(ix); return fail(PGI_ECUDA, std::string(e.what) + ": " + cudaGetErrorString(e.e)); }
      catch (const std::bad_alloc &) { pgi_index_free(ix); return fail(PGI_ENOMEM, "out of memory while building the index"); }
      catch (const std::exception &e) { pgi_index_free(ix); return fail(PGI_EINVAL, e.what()); }
}

// ------------------------------------------------------------------------------------------ cached device image
extern "C" int pgi_image_write(const void *ri_bytes, size_t ri_len, const void *tag_bytes, size_t tag_len, int tag_layout, const char *path) {
    if ((!ri_bytes && !tag_bytes) || !path) return fail(PGI_EINVAL, "null argument");
    try {
        RiImage ri; TagImage tg; IndexView v;
        flatten_files(ri_bytes, ri_len, tag_bytes, tag_len, tag_layout, ri, tg, v);
        write_image(v, path, ri_bytes ? ri_len : 0, ri_bytes ? edge_sum(ri_bytes, ri_len) : 0,
                    (tag_bytes && tag_len) ? tag_len : 0, (tag_bytes && tag_len) ? edge_sum(tag_bytes, tag_len) : 0);
        return PGI_OK;
    } PGI_CATCH_ALL
}

static void image_info_of(const MappedImage &im, pgi_image_info_t *f) {
    std::memset(f, 0, sizeof(*f));
    const ImageHeader &h = *im.h;
    f->file_bytes = h.file_bytes; f->n = h.meta.k.n; f->sigma = h.meta.k.sigma; f->n_dev_blocks = h.meta.k.n_blocks; f->bucket_shift = h.meta.k.bshift;
    f->has_ri = h.meta.has_ri; f->has_tags = h.meta.has_tags; f->tag_layout = h.meta.tag_layout; f->has_locate = h.meta.has_locate; f->is_encoded = h.meta.encoded;
    f->src_ri_bytes = h.src_ri_bytes; f->src_tag_bytes = h.src_tag_bytes; f->src_ri_sum = h.src_ri_sum; f->src_tag_sum = h.src_tag_sum;
    for (int s = 0; s < SEC_COUNT; ++s) f->payload_bytes += h.sec[s].bytes;
}

extern "C" int pgi_image_info(const char *path, int verify, pgi_image_info_t *out) {
    if (!path || !out) return fail(PGI_EINVAL, "null argument");
    try {
        MappedImage im; im.open(path, verify != 0);
        image_info_of(im, out);
        return PGI_OK;
    } PGI_CATCH_ALL
}

extern "C" uint64_t pgi_image_source_sum(const void *bytes, size_t len) { return bytes ? edge_sum(bytes, len) : 0; }

extern "C" int pgi_index_open_image(const char *path, int device, uint64_t chunk_bytes, int verify, pgi_index **out, pgi_image_info_t *info) {
    if (!path || !out) return fail(PGI_EINVAL, "null argument");
    if (int rc = check_device(device)) return rc;
    try {
        const char *env = std::getenv("PGI_IMAGE_VERIFY");
        MappedImage im; im.open(path, verify != 0 || (env && env[0] == '1'));
        ::madvise((void *) im.base, im.len, MADV_SEQUENTIAL);
        IndexView v; im.view(v);
        if (!chunk_bytes) chunk_bytes = 64ull << 20;
        if (chunk_bytes < 4096) chunk_bytes = 4096;
        uint64_t chunks = 0;
        *out = upload_view(v, device, (size_t) chunk_bytes, &chunks);
        if (info) { image_info_of(im, info); info->chunks_copied = chunks; info->chunk_bytes = chunk_bytes; }
        return PGI_OK;
    } PGI_CATCH_ALL
}

extern "C" int pgi_index_info(const pgi_index *ix, pgi_index_info_t *out) {
    if (!ix || !out) return fail(PGI_EINVAL, "null argument");
    *out = ix->info; return PGI_OK;
}

// ------------------------------------------------------------------------------------------ clone / free
template <typename T> static T *clone_arr(const T *src, int src_dev, uint64_t n, int dst_dev, std::vector<void *> &owned) {
    size_t bytes = (n ? n : 1) * sizeof(T);
    void *d = nullptr;
    CK(cudaMalloc(&d, bytes)); owned.push_back(d);
    if (n) CK(cudaMemcpyPeer(d, dst_dev, src, src_dev, n * sizeof(T)));     // device to device over NVLink, no host bounce
    else CK(cudaMemset(d, 0, bytes));
    return (T *) d;
}

extern "C" int pgi_index_clone_to(const pgi_index *src, int device, pgi_index **out) {
    if (!src || !out) return fail(PGI_EINVAL, "null argument");
    if (int rc = check_device(device)) return rc;
    if (src->host_resident_bytes) return fail(PGI_EINVAL, "an out-of-core index (run blocks in host memory) is shared, not cloned: open it once per GPU");
    pgi_index *ix = nullptr;
    try {
        DeviceGuard g(device);
        ix = new pgi_index(); ix->device = device; ix->consts = src->consts; ix->ivk = src->ivk; ix->info = src->info;
        cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, device)); ix->sm_count = prop.multiProcessorCount;
        ix->ri = src->ri; ix->tg = src->tg; ix->has_tags = src->has_tags; ix->has_ri = src->has_ri;
        ix->loc = src->loc; ix->has_locate = src->has_locate;
        ix->n_rs_dir = src->n_rs_dir; ix->n_rs_low = src->n_rs_low; ix->n_last_dir = src->n_last_dir; ix->n_last_low = src->n_last_low;
        ix->n_samples_words = src->n_samples_words; ix->n_ltr_words = src->n_ltr_words;
        ix->n_block_words = src->n_block_words; ix->n_quirk = src->n_quirk; ix->n_iv_dir = src->n_iv_dir;
        ix->n_iv_low = src->n_iv_low; ix->n_starts = src->n_starts; ix->n_words = src->n_words; ix->n_stream = src->n_stream;
        const int sd = src->device;
        ix->ri.blocks = clone_arr(src->ri.blocks, sd, src->n_block_words, device, ix->owned);
        ix->ri.q_start = clone_arr(src->ri.q_start, sd, src->n_quirk, device, ix->owned);
        ix->ri.q_val = clone_arr(src->ri.q_val, sd, src->n_quirk, device, ix->owned);
        ix->d_desc = clone_arr(src->d_desc, sd, PGI_DESC_WORDS, device, ix->owned);
        ix->d_desc_c = clone_arr(src->d_desc_c, sd, PGI_DESC_WORDS, device, ix->owned);
        ix->seed_k = src->seed_k;
        if (src->seed_k) ix->d_seed = clone_arr(src->d_seed, sd, (uint64_t) 1 << (2u * src->seed_k), device, ix->owned);
        if (src->has_locate) {
            ix->loc.run_starts.dir = clone_arr(src->loc.run_starts.dir, sd, src->n_rs_dir, device, ix->owned);
            ix->loc.run_starts.low = clone_arr(src->loc.run_starts.low, sd, src->n_rs_low, device, ix->owned);
            ix->loc.last.dir = clone_arr(src->loc.last.dir, sd, src->n_last_dir, device, ix->owned);
            ix->loc.last.low = clone_arr(src->loc.last.low, sd, src->n_last_low, device, ix->owned);
            ix->loc.samples.words = clone_arr(src->loc.samples.words, sd, src->n_samples_words, device, ix->owned);
            ix->loc.last_to_run.words = clone_arr(src->loc.last_to_run.words, sd, src->n_ltr_words, device, ix->owned);
        }
        if (src->has_tags) {
            ix->tg.iv.dir = clone_arr(src->tg.iv.dir, sd, src->n_iv_dir, device, ix->owned);
            ix->tg.iv.low = clone_arr(src->tg.iv.low, sd, src->n_iv_low, device, ix->owned);
            ix->tg.starts = clone_arr(src->tg.starts, sd, src->n_starts, device, ix->owned);
            ix->tg.words = clone_arr(src->tg.words, sd, src->n_words, device, ix->owned);
            ix->tg.stream = clone_arr(src->tg.stream, sd, src->n_stream, device, ix->owned);
        }
        CK(cudaDeviceSynchronize());
        *out = ix; return PGI_OK;
    } catch (const CudaError &e) { pgi_index_free(ix); return fail(PGI_ECUDA, std::string(e.what) + ": " + cudaGetErrorString(e.e)); }
}

// ------------------------------------------------------------------------------------------ batch context
enum { EV_START = 0, EV_MEM, EV_ORDER, EV_PLAN, EV_DECODE, EV_END, EV_K1A, EV_K1B, EV_COUNT };   // K1A / K1B: around the K1 launch alone

struct pgi_batch {
    pgi_index *ix = nullptr;
    cudaStream_t stream = nullptr; bool own_stream = false;
    cudaEvent_t ev[EV_COUNT]; bool ev_ok = false;
    // inputs
    DevBuf d_reads, d_read_off; const uint64_t *reads64 = nullptr; const uint64_t *read_off = nullptr; uint64_t n_reads = 0, reads_bytes = 0;
    DevBuf d_slot_ctr;                                       // partitioned MEM-slot counters (MemArgs::slot_ctr)
    DevBuf d_codes, d_read_flags;                            // 2-bit codes of the reads + per-read "other bytes" flags (seed lookups)
    // MEM stage
    DevBuf d_ctr, d_mem_buf, d_mem_key, d_nmems, d_mem_off, d_mems, d_cub;
    // tag stage
    DevBuf d_n_runs, d_cnt, d_raw_off, d_first, d_skip, d_raw, d_raw2, d_ucnt, d_tag_off, d_tags, d_large, d_seg_b, d_seg_e;
    PinBuf h_ctr, h_mem_off, h_mems, h_tag_off, h_tags, h_n_runs;
    DevBuf d_c_cnt, d_c_rec, d_c_rt, d_c_tags; PinBuf h_c_cnt, h_c_rec, h_c_rt, h_c_tags;   // compact wire form (pgi_batch_fetch_compact)
    double mems_per_read = 4.0, raw_per_mem = 2.0;         // running estimates that size the next step's buffers
    uint64_t mem_cap_hint = 0, raw_cap_hint = 0, large_cap_hint = 0;   // raised when a step did not fit (collect_and_check)
    uint64_t mems_cap = 0, mem_slots = 0, part_cap = 0, raw_cap = 0, large_cap = 0;   // capacities the current step was launched with
    uint64_t n_mems = 0, n_tags = 0, extensions = 0, seed_lookups = 0, blocks_fetched = 0;
    double stage_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    bool has_result = false;
    bool want_compact = false;                              // pack the result into the compact wire form INSIDE the step (pgi_find_mems_compact)
    bool compact_done = false;                              // ... and this step did
    bool chain_k1 = false;                                  // this step's K1 launch is ordered behind the previous step's (PGI_CHAIN=3)
    bool run_pending = false, run_count_mode = false, run_mutex_held = false; uint64_t run_min_len = 0, run_min_occ = 0; int run_launches = 0;   // step enqueued, not yet waited for
    void release_all() {
        DevBuf *db[] = {&d_slot_ctr, &d_codes, &d_read_flags, &d_reads, &d_read_off, &d_ctr, &d_mem_buf, &d_mem_key, &d_nmems, &d_mem_off, &d_mems, &d_cub, &d_n_runs, &d_cnt,
                        &d_raw_off, &d_first, &d_skip, &d_raw, &d_raw2, &d_ucnt, &d_tag_off, &d_tags, &d_large, &d_seg_b, &d_seg_e, &d_c_cnt, &d_c_rec, &d_c_rt, &d_c_tags};
        for (DevBuf *b : db) b->release();
        PinBuf *pb[] = {&h_ctr, &h_mem_off, &h_mems, &h_tag_off, &h_tags, &h_n_runs, &h_c_cnt, &h_c_rec, &h_c_rt, &h_c_tags};
        for (PinBuf *b : pb) b->release();
    }
};

extern "C" void pgi_batch_free(pgi_batch *b) {
    if (!b) return;
    int prev = -1; cudaGetDevice(&prev); cudaSetDevice(b->ix->device);
    if (b->stream) cudaStreamSynchronize(b->stream);
    b->release_all();
    if (b->ev_ok) for (int i = 0; i < EV_COUNT; ++i) cudaEventDestroy(b->ev[i]);
    if (b->own_stream && b->stream) cudaStreamDestroy(b->stream);
    if (prev >= 0) cudaSetDevice(prev);
    delete b;
}

extern "C" void pgi_index_free(pgi_index *ix) {
    if (!ix) return;
    for (pgi_batch *b : ix->pool) pgi_batch_free(b);
    int prev = -1; cudaGetDevice(&prev); cudaSetDevice(ix->device);
    if (ix->ev_end) cudaEventDestroy(ix->ev_end);
    if (ix->ev_k1) cudaEventDestroy(ix->ev_k1);
    for (void *p : ix->owned) cudaFree(p);
    for (void *p : ix->owned_host) cudaFreeHost(p);
    if (prev >= 0) cudaSetDevice(prev);
    delete ix;
}

extern "C" int pgi_batch_create(pgi_index *ix, void *cuda_stream, pgi_batch **out) {
    if (!ix || !out) return fail(PGI_EINVAL, "null argument");
    pgi_batch *b = nullptr;
    try {
        DeviceGuard g(ix->device);
        b = new pgi_batch(); b->ix = ix;
        if (cuda_stream) b->stream = (cudaStream_t) cuda_stream;
        else { CK(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking)); b->own_stream = true; }
        for (int i = 0; i < EV_COUNT; ++i) CK(cudaEventCreate(&b->ev[i]));
        b->ev_ok = true;
        b->d_ctr.need(256); b->h_ctr.need(256);
        *out = b; return PGI_OK;
    } catch (const CudaError &e) { if (b) { b->ix = ix; pgi_batch_free(b); } return fail(PGI_ECUDA, std::string(e.what) + ": " + cudaGetErrorString(e.e)); }
}

extern "C" int pgi_batch_upload(pgi_batch *b, const char *reads, const uint64_t *read_off, uint64_t n_reads) {
    if (!b || (!reads && n_reads) || !read_off) return fail(PGI_EINVAL, "null argument");
    try {
        DeviceGuard g(b->ix->device);
        if (read_off[0] != 0) return fail(PGI_EINVAL, "read_off[0] must be 0");
        for (uint64_t i = 0; i < n_reads; ++i) if (read_off[i + 1] < read_off[i]) return fail(PGI_EINVAL, "read offsets must be ascending");
        const uint64_t total = read_off[n_reads];
        b->d_reads.need(total + 16); b->d_read_off.need((n_reads + 1) * 8);
        if (total) CK(cudaMemcpyAsync(b->d_reads.p, reads, total, cudaMemcpyHostToDevice, b->stream));
        CK(cudaMemsetAsync((char *) b->d_reads.p + total, 0, 16, b->stream));
        CK(cudaMemcpyAsync(b->d_read_off.p, read_off, (n_reads + 1) * 8, cudaMemcpyHostToDevice, b->stream));
        b->reads64 = b->d_reads.as<uint64_t>(); b->read_off = b->d_read_off.as<uint64_t>(); b->n_reads = n_reads; b->has_result = false; b->compact_done = false;
        b->reads_bytes = total;
        return PGI_OK;
    } catch (const CudaError &e) { return fail(PGI_ECUDA, std::string(e.what) + ": " + cudaGetErrorString(e.e)); }
}

extern "C" int pgi_batch_set_device_reads(pgi_batch *b, const char *d_reads, const uint64_t *d_read_off, uint64_t n_reads, uint64_t total_bytes) {
    if (!b || !d_read_off || (!d_reads && n_reads)) return fail(PGI_EINVAL, "null argument");
    if (((uintptr_t) d_reads & 7) != 0) return fail(PGI_EINVAL, "device reads must be 8-byte aligned");
    b->reads_bytes = total_bytes;
    b->reads64 = (const uint64_t *) d_reads; b->read_off = d_read_off; b->n_reads = n_reads; b->has_result = false; b->compact_done = false;
    return PGI_OK;
}

// ------------------------------------------------------------------------------------------ the pipeline
namespace {

inline unsigned grid_for(uint64_t n, unsigned threads) { return (unsigned) std::max<uint64_t>(1, (n + threads - 1) / threads); }

// Exclusive prefix sum of n + 1 u64 values (the last input is forced to 0 by the caller) -> out[0..n].
void exclusive_scan(pgi_batch *b, const uint64_t *in, uint64_t *out, uint64_t n_plus_1, int &launches) {
    size_t tmp = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, in, out, (int64_t) n_plus_1, b->stream));
    b->d_cub.need(tmp);
    CK(cub::DeviceScan::ExclusiveSum(b->d_cub.p, tmp, in, out, (int64_t) n_plus_1, b->stream));
    (void) launches;                                  // library (CUB) launches are not counted as ours
}

// ---- One find_mems step is ONE stream-ordered sequence of launches with ONE host synchronisation at its end.
// Every size a later stage needs (MEMs found, items to decode, long lists, unique tags) stays on the device: kernels are
// launched over the CAPACITY of their buffers and read the count from the state words below; capacities come from the
// previous steps on the same batch context (grow-only, with head-room), and the rare step that does not fit is detected
// from the counters fetched at the end and run again with larger buffers. (Five host round trips per step -- sizing each
// stage from the previous one's count -- exposed every step to five scheduler wake-ups; on a box with 8 ranks that was the
// limiter of the device-resident rate: one rank at 15.6 ms per step beside seven at 9.5 ms, SCALE_r01.)
//
// device state words (d_ctr, u64): [0] read cursor  [1] MEM slots taken (single counter)  [2] FMD extensions  [4] seed lookups
//   [8] MEMs found (exact even after an overflow)   [9] slots the fullest partition asked for   [10] MEMs the later stages
//   see (= [8] clamped to the capacity of the ordered MEM buffer)   [12] long tag lists queued   [13] raw buffer too small
// host mirror (h_ctr, pinned; written by collect_kernel): [0] = [8], [1] = [9], [2] extensions, [3] seed lookups,
//   [4] items to decode (raw_off[nm]), [5] long lists, [6] raw overflow flag, [7] unique tags (tag_off[nm]), [9] run blocks fetched
// A counter reaches the host through a KERNEL writing pinned memory (unified addressing), not a cudaMemcpyAsync: a copy,
// however small, queues on the device-to-host copy engine behind the result download of whichever call finished before.
enum { ST_CURSOR = 0, ST_SLOTS = 1, ST_EXT = 2, ST_SEED = 4, ST_BLK = 5, ST_BADOFF = 6, ST_NMEMS = 8, ST_MAXPART = 9, ST_NM = 10, ST_LARGE = 12, ST_RAWOVF = 13, ST_NOFIT = 14, ST_WORDS = 16 };

__global__ void mem_finalize_kernel(unsigned long long *st, const unsigned long long *slot_ctr, uint64_t part_cap, uint64_t mems_cap) {
    __shared__ unsigned long long s_sum[PGI_SLOT_PARTS], s_max[PGI_SLOT_PARTS];
    const unsigned t = threadIdx.x;
    s_sum[t] = s_max[t] = part_cap ? slot_ctr[16u * t] : 0ull;
    __syncthreads();
    for (unsigned o = PGI_SLOT_PARTS / 2; o; o >>= 1) {
        if (t < o) { s_sum[t] += s_sum[t + o]; if (s_max[t + o] > s_max[t]) s_max[t] = s_max[t + o]; }
        __syncthreads();
    }
    if (t == 0) {
        const unsigned long long total = part_cap ? s_sum[0] : st[ST_SLOTS];
        st[ST_NMEMS] = total; st[ST_MAXPART] = part_cap ? s_max[0] : total;
        st[ST_NM] = total < mems_cap ? total : mems_cap;
    }
}
__global__ void set_nm_kernel(unsigned long long *st, uint64_t nm) { st[ST_NMEMS] = nm; st[ST_MAXPART] = 0; st[ST_NM] = nm; }
__global__ void collect_kernel(const unsigned long long *st, const uint64_t *raw_off, const uint64_t *tag_off, volatile uint64_t *host) {
    if (threadIdx.x == 0) {
        const uint64_t nm = st[ST_NM];
        host[0] = st[ST_NMEMS]; host[1] = st[ST_MAXPART]; host[2] = st[ST_EXT]; host[3] = st[ST_SEED];
        host[4] = raw_off ? raw_off[nm] : 0; host[5] = st[ST_LARGE]; host[6] = st[ST_RAWOVF]; host[7] = tag_off ? tag_off[nm] : 0; host[9] = st[ST_BLK]; host[10] = st[ST_BADOFF]; host[12] = st[ST_NOFIT];
        __threadfence_system();
    }
}
__global__ void peek_kernel(const uint64_t *src, volatile uint64_t *host_dst, unsigned n) {
    if (threadIdx.x < n) host_dst[threadIdx.x] = src[threadIdx.x];
    __threadfence_system();
}
uint64_t read_back(pgi_batch *b, const void *dptr) {              // locate only (not on the find_mems path)
    peek_kernel<<<1, 32, 0, b->stream>>>((const uint64_t *) dptr, b->h_ctr.as<uint64_t>() + 8, 1);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(b->stream));
    return b->h_ctr.as<volatile uint64_t>()[8];
}

// K1 instantiations: block-fetch mode x CTA shape (pgi_kernels.cuh). PGI_K1=<name> picks one for tuning runs
// (e.g. PGI_K1=04 is the lane-private fetch, 256 threads x 4 CTAs per SM); the default is the measured best.
// Default: mode 4 (padded staging + L2 evict-first hint) for an index whose run blocks live in DRAM -- config 4: 7.62 ms
// against 7.79 ms for mode 1 -- but mode 1 for one whose run blocks fit the L2: there the kernel is bound by issue / shared
// memory, not by DRAM, and the padded staging costs 2.3 x (shipped xy index, 1 M reads: 2.79 ms against 6.45 ms, profiles/r03o).
#define PGI_K1_HOT_INDEX_BYTES (64ull << 20)
struct K1Variant { const char *name; void (*fn)(const MemArgs); unsigned threads; };
const K1Variant &k1_variant(const pgi_index *ix) {
    static const K1Variant table[] = {          // first = default for DRAM-resident indexes (tools/k1_sweep.py, tools/index_sweep.py)
        {"44", fmd_mem_kernel<4, 256, 4>, 256}, {"24", fmd_mem_kernel<2, 256, 4>, 256}, {"14", fmd_mem_kernel<1, 256, 4>, 256}, {"1a", fmd_mem_kernel<1, 128, 8>, 128}, {"1b", fmd_mem_kernel<1, 128, 9>, 128},
        {"15", fmd_mem_kernel<1, 256, 5>, 256}, {"04", fmd_mem_kernel<0, 256, 4>, 256}, {"2a", fmd_mem_kernel<2, 128, 8>, 128},
    };
    const char *e = std::getenv("PGI_K1");                  // read per call: tools/k1_sweep.py switches variants in one process
    if (e) for (const K1Variant &v : table) if (!std::strcmp(v.name, e)) return v;
    return ix->n_block_words * 8ull <= PGI_K1_HOT_INDEX_BYTES ? table[2] : table[0];
}

// The persistent grid is sized for the occupancy the kernel's static shared memory allows (4 CTAs x 43 KB per SM). With the
// default carve-out preference the driver chooses the shared-memory configuration itself; asking for the smallest one that
// holds all those CTAs makes the occupancy the grid was sized for a fact instead of a heuristic.
static void prefer_max_shared(void (*fn)(const MemArgs)) {
    const void *f = reinterpret_cast<const void *>(fn);
    cudaFuncAttributes at; CK(cudaFuncGetAttributes(&at, f));
    int dev = 0, smem_sm = 0, blocks = 0; CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev));
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, f, at.maxThreadsPerBlock, 0));
    int pct = (int) ((100.0 * blocks * (double) (at.sharedSizeBytes + 1024)) / (double) smem_sm) + 1;
    if (const char *e = std::getenv("PGI_CARVEOUT")) pct = std::atoi(e);
    if (pct > 100) pct = 100;
    CK(cudaFuncSetAttribute(f, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
}
static unsigned persistent_grid(pgi_index *ix, void (*fn)(const MemArgs), unsigned threads, uint64_t n) {
    // the occupancy query and the carve-out are per function: asked once, not in every step
    static std::mutex mu; static std::vector<std::pair<const void *, int>> cache;
    int bps = 0;
    { std::lock_guard<std::mutex> l(mu); for (auto &c : cache) if (c.first == (const void *) fn) bps = c.second; }
    if (!bps) {
        prefer_max_shared(fn);
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, fn, (int) threads, 0));
        if (bps < 1) bps = 1;
        std::lock_guard<std::mutex> l(mu); cache.emplace_back((const void *) fn, bps);
    }
    return (unsigned) std::max<uint64_t>(1, std::min<uint64_t>((uint64_t) ix->sm_count * bps, (n + threads - 1) / threads));
}

// MEM finding: K0 (2-bit codes), K1, the count of MEMs (device), scan of the per-read counts, scatter into read order.
// Enqueues only. Leaves b->mems_cap = capacity of the ordered MEM buffer the later stages are launched over.
void enqueue_mem_stage(pgi_batch *b, uint64_t min_len, uint64_t min_occ, int &launches) {
    pgi_index *ix = b->ix; const uint64_t n = b->n_reads;
    b->d_nmems.need((n + 1) * 8); b->d_mem_off.need((n + 1) * 8);
    const uint64_t cap = std::max<uint64_t>(1024, b->mem_cap_hint ? b->mem_cap_hint : (uint64_t) ((double) n * b->mems_per_read * 1.25) + 1024);
    const K1Variant &k1 = k1_variant(ix);
    const unsigned grid = persistent_grid(ix, k1.fn, k1.threads, n);
    unsigned long long *st = b->d_ctr.as<unsigned long long>();
    if (ix->seed_k && n) {                                   // K0: 2-bit codes + per-read flags for the seed lookups
        const uint64_t n_words = (b->reads_bytes + 7) / 8;
        b->d_codes.need(n_words * 2 + 16); b->d_read_flags.need((n / 32 + 1) * 4);
        CK(cudaMemsetAsync(b->d_read_flags.p, 0, (n / 32 + 1) * 4, b->stream));
        CK(cudaMemsetAsync((char *) b->d_codes.p + n_words * 2, 0, 16, b->stream));
        if (n_words) {
            pack_reads_kernel<<<grid_for(n_words, 256), 256, 0, b->stream>>>(b->reads64, n_words, b->reads_bytes, b->read_off, n,
                                                                              b->d_codes.as<uint16_t>(), b->d_read_flags.as<uint32_t>());
            ++launches; CK(cudaGetLastError());
        }
    }
    // MEM slots: one global counter, or -- for batches of PGI_SLOT_PARTS=<reads> (default 600000; 0 = never; 1 = always)
    // reads and more -- 256 partitions with a counter each (pgi_kernels.cuh MemArgs)
    uint64_t parts_min = 600000;
    if (const char *e = std::getenv("PGI_SLOT_PARTS")) parts_min = std::strtoull(e, nullptr, 10);
    const bool parts_on = parts_min != 0 && n >= parts_min;
    const uint64_t part_cap = parts_on ? (cap + PGI_SLOT_PARTS - 1) / PGI_SLOT_PARTS + 64 : 0;
    const uint64_t slots = parts_on ? part_cap * PGI_SLOT_PARTS : cap;
    b->part_cap = part_cap; b->mem_slots = slots; b->mems_cap = slots;
    b->d_slot_ctr.need(PGI_SLOT_PARTS * 128);
    b->d_mem_buf.need(slots * 32); b->d_mem_key.need(slots * 8); b->d_mems.need((slots + 1) * 32);
    CK(cudaMemsetAsync(st, 0, ST_WORDS * 8, b->stream));
    CK(cudaMemsetAsync(b->d_slot_ctr.p, 0, PGI_SLOT_PARTS * 128, b->stream));
    CK(cudaMemsetAsync(b->d_nmems.p, 0, (n + 1) * 8, b->stream));
    MemArgs a;
    a.ix = ix->ri; a.desc = ix->d_desc; a.reads64 = b->reads64; a.read_off = b->read_off;
    a.n_reads = n; a.min_len = std::min<uint64_t>(min_len, 0x7FFFFFFFull); a.min_occ = min_occ;
    a.seed = ix->d_seed; a.seed_k = ix->seed_k; a.pad = 0; a.codes32 = b->d_codes.as<uint32_t>(); a.read_flags = b->d_read_flags.as<uint32_t>();
    a.cursor = st + ST_CURSOR; a.mem_count = st + ST_SLOTS; a.ext_count = st + ST_EXT; a.seed_count = st + ST_SEED; a.blk_count = st + ST_BLK; a.bad_off = st + ST_BADOFF; a.reads_bytes = b->reads_bytes; a.mem_cap = cap;
    a.slot_ctr = b->d_slot_ctr.as<unsigned long long>(); a.part_cap = part_cap;
    a.mem_buf = b->d_mem_buf.as<uint4>(); a.mem_key = b->d_mem_key.as<uint64_t>(); a.read_nmems = b->d_nmems.as<uint64_t>();
    if (b->chain_k1 && ix->ev_k1_set) CK(cudaStreamWaitEvent(b->stream, ix->ev_k1, 0));
    CK(cudaEventRecord(b->ev[EV_K1A], b->stream));
    if (n) { k1.fn<<<grid, k1.threads, 0, b->stream>>>(a); ++launches; CK(cudaGetLastError()); }
    CK(cudaEventRecord(b->ev[EV_K1B], b->stream));
    if (b->chain_k1) { CK(cudaEventRecord(ix->ev_k1, b->stream)); ix->ev_k1_set = true; }
    mem_finalize_kernel<<<1, PGI_SLOT_PARTS, 0, b->stream>>>(st, a.slot_ctr, part_cap, b->mems_cap);
    ++launches; CK(cudaGetLastError());
    CK(cudaEventRecord(b->ev[EV_MEM], b->stream));
    exclusive_scan(b, b->d_nmems.as<uint64_t>(), b->d_mem_off.as<uint64_t>(), n + 1, launches);
    mem_scatter_kernel<<<grid_for(slots, 256), 256, 0, b->stream>>>(b->d_mem_buf.as<uint4>(), b->d_mem_key.as<uint64_t>(), b->d_mem_off.as<uint64_t>(), slots,
                                                                      b->d_mems.as<uint4>(), b->d_slot_ctr.as<unsigned long long>(), part_cap,
                                                                      (const uint64_t *) (st + ST_NMEMS), b->mems_cap);
    ++launches; CK(cudaGetLastError());
    CK(cudaEventRecord(b->ev[EV_ORDER], b->stream));
}

// Whole-read backward search (query_tags path): the same persistent kernel in OP 1; one record per read, in read order.
__global__ void iota_kernel(uint64_t n_plus_1, uint64_t *out) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i < n_plus_1) out[i] = i;
}
void enqueue_count_stage(pgi_batch *b, int &launches) {
    pgi_index *ix = b->ix; const uint64_t n = b->n_reads;
    b->d_mem_off.need((n + 1) * 8); b->d_mems.need((n + 1) * 32);
    // whole-read backward search (OP 1). Fetch mode 1 is the measured best HERE (4.37 ms per 2 M reads against 5.2 ms for
    // modes 2 and 4, profiles/r03k: nearly every LF step of a whole-read search ranks two positions in different buckets);
    // PGI_K1C = 2 / 4 selects the others for tuning runs
    void (*fn)(const MemArgs) = fmd_mem_kernel<1, 256, 4, 1>; const unsigned threads = 256;
    if (const char *e = std::getenv("PGI_K1C")) { if (e[0] == '4') fn = fmd_mem_kernel<4, 256, 4, 1>; else if (e[0] == '2') fn = fmd_mem_kernel<2, 256, 4, 1>; }
    const unsigned grid = persistent_grid(ix, fn, threads, n);
    unsigned long long *st = b->d_ctr.as<unsigned long long>();
    CK(cudaMemsetAsync(st, 0, ST_WORDS * 8, b->stream));
    MemArgs a;
    a.ix = ix->ri; a.desc = ix->d_desc_c; a.reads64 = b->reads64; a.read_off = b->read_off;
    a.n_reads = n; a.min_len = 1; a.min_occ = 0; a.seed = nullptr; a.seed_k = 0; a.pad = 0; a.codes32 = nullptr; a.read_flags = nullptr;
    a.cursor = st + ST_CURSOR; a.mem_count = st + ST_SLOTS; a.ext_count = st + ST_EXT; a.seed_count = st + ST_SEED; a.blk_count = st + ST_BLK; a.bad_off = st + ST_BADOFF; a.reads_bytes = b->reads_bytes; a.mem_cap = n; a.slot_ctr = nullptr; a.part_cap = 0;
    a.mem_buf = b->d_mems.as<uint4>(); a.mem_key = nullptr; a.read_nmems = nullptr;
    if (b->chain_k1 && ix->ev_k1_set) CK(cudaStreamWaitEvent(b->stream, ix->ev_k1, 0));
    CK(cudaEventRecord(b->ev[EV_K1A], b->stream));
    if (n) { fn<<<grid, threads, 0, b->stream>>>(a); ++launches; CK(cudaGetLastError()); }
    CK(cudaEventRecord(b->ev[EV_K1B], b->stream));
    if (b->chain_k1) { CK(cudaEventRecord(ix->ev_k1, b->stream)); ix->ev_k1_set = true; }
    iota_kernel<<<grid_for(n + 1, 256), 256, 0, b->stream>>>(n + 1, b->d_mem_off.as<uint64_t>());
    set_nm_kernel<<<1, 1, 0, b->stream>>>(st, n);
    launches += 2; CK(cudaGetLastError());
    b->mems_cap = n; b->part_cap = 0; b->mem_slots = n;
    CK(cudaEventRecord(b->ev[EV_MEM], b->stream)); CK(cudaEventRecord(b->ev[EV_ORDER], b->stream));
}

// Tag lookup for up to `cap` intervals (their number is st[ST_NM]) given either as ordered MEM records or as start/end
// arrays (device pointers). Enqueues only.
void enqueue_tag_stage(pgi_batch *b, uint64_t cap, const uint4 *mems, const uint64_t *start, const uint64_t *end, int &launches) {
    pgi_index *ix = b->ix; cudaStream_t s = b->stream;
    unsigned long long *st = b->d_ctr.as<unsigned long long>();
    const uint64_t *nm_dev = (const uint64_t *) (st + ST_NM);
    b->d_n_runs.need((cap + 1) * 8); b->d_cnt.need((cap + 1) * 8); b->d_raw_off.need((cap + 1) * 8); b->d_first.need((cap + 1) * 8);
    b->d_skip.need((cap + 1) * 4); b->d_ucnt.need((cap + 1) * 8); b->d_tag_off.need((cap + 1) * 8);
    const uint64_t raw_cap = std::max<uint64_t>(b->raw_cap_hint, (uint64_t) ((double) cap * std::max(1.0, b->raw_per_mem) * 1.25) + 4096);
    const uint64_t large_cap = b->large_cap_hint;
    b->raw_cap = raw_cap; b->large_cap = large_cap;
    b->d_raw.need(raw_cap * 8); b->d_tags.need(raw_cap * 8); b->d_large.need((large_cap + 1) * 8);
    CK(cudaMemsetAsync(b->d_cnt.p, 0, (cap + 1) * 8, s)); CK(cudaMemsetAsync(b->d_ucnt.p, 0, (cap + 1) * 8, s));
    CK(cudaMemsetAsync(b->d_n_runs.p, 0, (cap + 1) * 8, s));
    if (cap) {
        tag_plan_kernel<<<grid_for(cap, 256), 256, 0, s>>>(ix->tg, cap, nm_dev, mems, start, end, b->d_n_runs.as<uint64_t>(), b->d_cnt.as<uint64_t>(),
                                                           b->d_first.as<uint64_t>(), b->d_skip.as<uint32_t>());
        ++launches; CK(cudaGetLastError());
    }
    exclusive_scan(b, b->d_cnt.as<uint64_t>(), b->d_raw_off.as<uint64_t>(), cap + 1, launches);
    CK(cudaEventRecord(b->ev[EV_PLAN], s));
    if (cap) {
        tag_small_kernel<<<grid_for(cap, 128), 128, 0, s>>>(ix->tg, cap, nm_dev, b->d_cnt.as<uint64_t>(), b->d_raw_off.as<uint64_t>(), b->d_first.as<uint64_t>(),
                                                            b->d_skip.as<uint32_t>(), b->d_raw.as<uint64_t>(), raw_cap, b->d_ucnt.as<uint64_t>(),
                                                            b->d_large.as<uint64_t>(), large_cap, st + ST_LARGE);
        ++launches; CK(cudaGetLastError());
    }
    const unsigned lgrid = (unsigned) std::min<uint64_t>((uint64_t) ix->sm_count * 8, std::max<uint64_t>(1, (large_cap * 32 + 255) / 256));
    if (large_cap) {                                    // long lists (more than PGI_TAG_SMALL runs): warp per list + CUB segmented sort
        b->d_raw2.need(raw_cap * 8); b->d_seg_b.need(large_cap * 8); b->d_seg_e.need(large_cap * 8);
        CK(cudaMemsetAsync(b->d_seg_b.p, 0, large_cap * 8, s)); CK(cudaMemsetAsync(b->d_seg_e.p, 0, large_cap * 8, s));
        tag_large_decode_kernel<<<lgrid, 256, 0, s>>>(ix->tg, st + ST_LARGE, large_cap, b->d_large.as<uint64_t>(), b->d_cnt.as<uint64_t>(), b->d_raw_off.as<uint64_t>(),
                                                      b->d_first.as<uint64_t>(), b->d_skip.as<uint32_t>(), b->d_raw.as<uint64_t>(),
                                                      b->d_seg_b.as<uint64_t>(), b->d_seg_e.as<uint64_t>());
        CK(cudaGetLastError());
        size_t tmp = 0;
        CK(cub::DeviceSegmentedSort::SortKeys(nullptr, tmp, b->d_raw.as<uint64_t>(), b->d_raw2.as<uint64_t>(), (int64_t) raw_cap, (int64_t) large_cap,
                                              b->d_seg_b.as<int64_t>(), b->d_seg_e.as<int64_t>(), s));
        b->d_cub.need(tmp);
        CK(cub::DeviceSegmentedSort::SortKeys(b->d_cub.p, tmp, b->d_raw.as<uint64_t>(), b->d_raw2.as<uint64_t>(), (int64_t) raw_cap, (int64_t) large_cap,
                                              b->d_seg_b.as<int64_t>(), b->d_seg_e.as<int64_t>(), s));
        tag_large_unique_kernel<<<lgrid, 256, 0, s>>>(st + ST_LARGE, large_cap, b->d_large.as<uint64_t>(), b->d_cnt.as<uint64_t>(), b->d_raw_off.as<uint64_t>(),
                                                      b->d_raw2.as<uint64_t>(), b->d_raw.as<uint64_t>(), b->d_ucnt.as<uint64_t>());
        CK(cudaGetLastError());
        launches += 2;                                   // decode + unique (the segmented sort is CUB's)
    }
    CK(cudaEventRecord(b->ev[EV_DECODE], s));
    exclusive_scan(b, b->d_ucnt.as<uint64_t>(), b->d_tag_off.as<uint64_t>(), cap + 1, launches);
    if (cap) {
        tag_copy_kernel<<<grid_for(cap, 256), 256, 0, s>>>(cap, nm_dev, b->d_cnt.as<uint64_t>(), b->d_raw_off.as<uint64_t>(), b->d_tag_off.as<uint64_t>(),
                                                           b->d_raw.as<uint64_t>(), b->d_tags.as<uint64_t>());
        ++launches; CK(cudaGetLastError());
    }
    if (large_cap) {
        tag_large_copy_kernel<<<lgrid, 256, 0, s>>>(st + ST_LARGE, large_cap, b->d_large.as<uint64_t>(), b->d_raw_off.as<uint64_t>(),
                                                    b->d_tag_off.as<uint64_t>(), b->d_raw.as<uint64_t>(), b->d_tags.as<uint64_t>());
        ++launches; CK(cudaGetLastError());
    }
}

// ---- compact wire form of the result (pgi_result_compact)
__global__ void iota_mul_kernel(uint64_t n_plus_1, uint64_t mul, uint64_t *out) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i < n_plus_1) out[i] = i * mul;
}
// flags[0] set when a value does not fit the compact form
__global__ void compact_reads_kernel(uint64_t n_reads, const uint64_t *mem_off, uint32_t *cnt, unsigned long long *flags) {
    const uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (i >= n_reads) return;
    const uint64_t c = mem_off[i + 1] - mem_off[i];
    if (c >> 32) flags[0] = 1ull;
    cnt[i] = (uint32_t) c;
}
// n_mems = the count, or -- nm_dev != nullptr, launched inside the step -- the capacity, the count being read from device memory
__global__ void compact_mems_kernel(uint64_t n_mems, const uint64_t *nm_dev, const uint4 *mems, const uint64_t *tag_off, const uint64_t *n_runs,
                                    uint64_t *rec, uint32_t *runs_tags, unsigned long long *flags) {
    const uint64_t m = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (m >= n_mems || (nm_dev && m >= *nm_dev)) return;
    const uint4 a = mems[2 * m], b = mems[2 * m + 1];
    const uint64_t start = a.x | (uint64_t) a.y << 32, end = a.z | (uint64_t) a.w << 32, k = b.x | (uint64_t) b.y << 32, sz = b.z | (uint64_t) b.w << 32;
    const uint64_t t0 = tag_off[m], nt = tag_off[m + 1] - t0, nr = n_runs[m];
    if ((start >> 24) || ((end - start) >> 24) || (k >> 40) || (sz >> 40) || (nt >> 32) || (nr >> 32)) flags[0] = 1ull;
    rec[2 * m] = (k & 0xFFFFFFFFFFull) | start << 40; rec[2 * m + 1] = (sz & 0xFFFFFFFFFFull) | (end - start) << 40;
    runs_tags[2 * m] = (uint32_t) nr; runs_tags[2 * m + 1] = (uint32_t) nt;
}
// The tags of the batch narrowed to tag_bytes (5 or 8) each: one thread per FOUR tags, which are 20 bytes = five aligned
// 32-bit words in the 5-byte form (the per-MEM loop of byte stores this replaces took a third of the packing time).
// n_tags = tag_off[nm], nm read from device memory inside a step (nm_dev) or given (nm_host).
__global__ void compact_tags_kernel(uint64_t cap_groups, const uint64_t *nm_dev, uint64_t nm_host, const uint64_t *tag_off, const uint64_t *tags, unsigned tag_bytes,
                                    uint32_t *ctags32, unsigned long long *flags) {
    const uint64_t g = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x;
    if (g >= cap_groups) return;
    const uint64_t nt = tag_off[nm_dev ? *nm_dev : nm_host], first = 4 * g;
    if (first >= nt) return;
    uint64_t v[4];
    for (unsigned i = 0; i < 4; ++i) v[i] = first + i < nt ? tags[first + i] : 0ull;
    if (tag_bytes == 8) { for (unsigned i = 0; i < 4; ++i) if (first + i < nt) reinterpret_cast<uint64_t *>(ctags32)[first + i] = v[i]; return; }
    if ((v[0] | v[1] | v[2] | v[3]) >> 40) flags[0] = 1ull;
    uint32_t *d = ctags32 + 5 * g;
    d[0] = (uint32_t) v[0];
    d[1] = (uint32_t) ((v[0] >> 32) & 0xFFu) | (uint32_t) (v[1] & 0xFFFFFFu) << 8;
    d[2] = (uint32_t) ((v[1] >> 24) & 0xFFFFu) | (uint32_t) (v[2] & 0xFFFFu) << 16;
    d[3] = (uint32_t) ((v[2] >> 16) & 0xFFFFFFu) | (uint32_t) (v[3] & 0xFFu) << 24;
    d[4] = (uint32_t) (v[3] >> 8);
}

unsigned compact_tag_bytes(const pgi_index *ix) { return (ix->has_tags && ix->tg.layout == 0 && ix->tg.width <= 40) ? 5u : 8u; }
// The compact wire form packed INSIDE the step (behind the tag stage, over the step's capacities, counts read from device
// memory), so that the download of a finished step is copies only. Packed behind the step instead, the two kernels wait for
// the NEXT caller's persistent fmd_mem_kernel to leave the SMs (9 ms) before the download can even start.
void enqueue_compact(pgi_batch *b, bool with_tags, int &launches) {
    const uint64_t n = b->n_reads, cap = b->mems_cap; cudaStream_t st = b->stream;
    const unsigned tag_bytes = compact_tag_bytes(b->ix);
    unsigned long long *stw = b->d_ctr.as<unsigned long long>();
    b->d_c_cnt.need((n + 1) * 4); b->d_c_rec.need((cap + 1) * 16); b->d_c_rt.need((cap + 1) * 8);
    const uint64_t tag_cap = with_tags ? b->raw_cap : 0;
    b->d_c_tags.need((tag_cap + 1) * (uint64_t) tag_bytes + 32);
    if (n) { compact_reads_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, b->d_mem_off.as<uint64_t>(), b->d_c_cnt.as<uint32_t>(), stw + ST_NOFIT); ++launches; }
    if (cap) {
        compact_mems_kernel<<<grid_for(cap, 256), 256, 0, st>>>(cap, (const uint64_t *) (stw + ST_NM), b->d_mems.as<uint4>(), b->d_tag_off.as<uint64_t>(), b->d_n_runs.as<uint64_t>(),
                                                                b->d_c_rec.as<uint64_t>(), b->d_c_rt.as<uint32_t>(), stw + ST_NOFIT);
        ++launches;
    }
    if (tag_cap) {
        compact_tags_kernel<<<grid_for(tag_cap / 4 + 1, 256), 256, 0, st>>>(tag_cap / 4 + 1, (const uint64_t *) (stw + ST_NM), 0, b->d_tag_off.as<uint64_t>(), b->d_tags.as<uint64_t>(), tag_bytes,
                                                                            b->d_c_tags.as<uint32_t>(), stw + ST_NOFIT);
        ++launches;
    }
    CK(cudaGetLastError());
}

// The ONE synchronisation of a step: fetch the counters, decide whether everything fitted. Returns true when the step has
// to be run again (capacity hints of the batch were raised).
void enqueue_collect(pgi_batch *b, bool with_tags) {
    unsigned long long *st = b->d_ctr.as<unsigned long long>();
    collect_kernel<<<1, 32, 0, b->stream>>>(st, with_tags ? b->d_raw_off.as<uint64_t>() : nullptr, with_tags ? b->d_tag_off.as<uint64_t>() : nullptr, b->h_ctr.as<uint64_t>());
    CK(cudaGetLastError());
    CK(cudaEventRecord(b->ev[EV_END], b->stream));
}
bool wait_and_check(pgi_batch *b, bool with_tags, bool mem_stage) {
    CK(cudaStreamSynchronize(b->stream));
    volatile uint64_t *h = b->h_ctr.as<volatile uint64_t>();
    const uint64_t total = h[0], maxpart = h[1], total_raw = h[4], n_large = h[5], raw_ovf = h[6];
    if (h[10]) throw std::invalid_argument("read offsets are not ascending or point beyond the reads buffer");
    b->extensions = h[2]; b->seed_lookups = h[3]; b->blocks_fetched = h[9]; b->n_mems = total; b->n_tags = with_tags ? h[7] : 0;
    bool again = false;
    if (mem_stage) {
        if (b->part_cap ? maxpart > b->part_cap : total > b->mem_slots) {
            // the counts are exact: size for them (partitioned: every partition gets what the fullest one asked for)
            b->mem_cap_hint = b->part_cap ? (maxpart + maxpart / 8 + 64) * PGI_SLOT_PARTS : total + total / 8 + 1024;
            again = true;
        } else if (b->n_reads) { b->mems_per_read = std::max(1.0, (double) total / (double) b->n_reads); b->mem_cap_hint = 0; }
    }
    if (with_tags && !again) {
        if (raw_ovf || total_raw > b->raw_cap) { b->raw_cap_hint = std::max<uint64_t>(2 * b->raw_cap, total_raw + total_raw / 4 + 4096); again = true; }
        if (n_large > b->large_cap) { b->large_cap_hint = std::max<uint64_t>(1024, n_large + n_large / 2); again = true; }
        if (!again && total) b->raw_per_mem = std::max(1.0, (double) total_raw / (double) total);
    }
    return again;
}

float ev_ms(pgi_batch *b, int from, int to) { float ms = 0; cudaEventElapsedTime(&ms, b->ev[from], b->ev[to]); return ms; }

} // namespace

#define PGI_CATCH \
    catch (const CudaError &e) { return fail(PGI_ECUDA, std::string(e.what) + ": " + cudaGetErrorString(e.e)); } \
    catch (const std::bad_alloc &) { return fail(PGI_ENOMEM, "out of host memory"); } \
    catch (const std::exception &e) { return fail(PGI_EINVAL, e.what()); }

// A step = batch_enqueue (everything of the step queued on the batch's stream, nothing waited for) + batch_finish (the one
// synchronisation: counters fetched, the rare step that did not fit its buffers run again).
// Steps enqueued on one index by concurrent callers or by several batches in flight are ordered ON THE GPU (PGI_CHAIN, read
// per step; the mutex is held while a step is enqueued, never while one runs):
//   1  a step starts when the step enqueued before it has ended (chain event ev_end): one step at a time, but the next one is
//      already queued when this one ends -- no host wake-up, mutex hand-over and two dozen launches between two steps
//   3  only the fmd_mem_kernel launches are ordered (ev_k1): the next step's persistent grid starts when this one's has
//      ended, and this step's short ordering / tag / packing kernels run beside it
//   2  no ordering: the streams overlap as the hardware schedules them
//   0  (one-call API only) the round-1 scheme: the caller holds the mutex until its step has finished
static int chain_mode() {
    if (const char *e = std::getenv("PGI_CHAIN")) { const int m = std::atoi(e); if (m >= 0 && m <= 3) return m; }
    return 1;
}
static int batch_enqueue(pgi_batch *b, bool count_mode, uint64_t min_len, uint64_t min_occ, bool mutex_held = false) {
    if (!b) return fail(PGI_EINVAL, "null argument");
    if (!b->read_off) return fail(PGI_EINVAL, "no reads staged: call pgi_batch_upload or pgi_batch_set_device_reads first");
    if (b->n_reads >> 32) return fail(PGI_EINVAL, "more than 2^32 reads in one batch");
    if (!b->ix->has_ri) return fail(PGI_EINVAL, "index was created without an r-index (tags only)");
    if (count_mode && !b->ix->consts.count_ok)
        return fail(PGI_EFORMAT, "encoded r-index without N: the reference's rankAt_encoded reads a run header as a sixth cumulative on such a "
                                 "file (src/r-index.cpp:577), its count_encoded results are not reproduced");
    try {
        DeviceGuard g(b->ix->device);
        const bool with_tags = b->ix->has_tags;
        int launches = 0;
        b->has_result = false; b->compact_done = false;
        b->run_count_mode = count_mode; b->run_min_len = min_len; b->run_min_occ = min_occ; b->run_mutex_held = mutex_held;
        pgi_index *ix = b->ix;
        const int mode = mutex_held ? 0 : chain_mode();
        std::unique_lock<std::mutex> order(ix->compute_mu, std::defer_lock);
        if (!mutex_held && mode != 2) order.lock();
        if ((mode == 1 || mode == 3) && !ix->ev_end) {
            CK(cudaEventCreateWithFlags(&ix->ev_end, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&ix->ev_k1, cudaEventDisableTiming));
        }
        b->chain_k1 = mode == 3;
        if (mode == 1 && ix->ev_end_set) CK(cudaStreamWaitEvent(b->stream, ix->ev_end, 0));
        CK(cudaEventRecord(b->ev[EV_START], b->stream));
        if (count_mode) enqueue_count_stage(b, launches);
        else enqueue_mem_stage(b, min_len, min_occ, launches);
        if (with_tags) enqueue_tag_stage(b, b->mems_cap, b->d_mems.as<uint4>(), nullptr, nullptr, launches);
        else {
            b->d_tag_off.need((b->mems_cap + 1) * 8); b->d_n_runs.need((b->mems_cap + 1) * 8); b->d_tags.need(8);
            CK(cudaMemsetAsync(b->d_tag_off.p, 0, (b->mems_cap + 1) * 8, b->stream));
            CK(cudaMemsetAsync(b->d_n_runs.p, 0, (b->mems_cap + 1) * 8, b->stream));
            CK(cudaEventRecord(b->ev[EV_PLAN], b->stream)); CK(cudaEventRecord(b->ev[EV_DECODE], b->stream));
        }
        if (b->want_compact) enqueue_compact(b, with_tags, launches);
        ++launches;                                          // collect_kernel
        enqueue_collect(b, with_tags);
        if (mode == 1) { CK(cudaEventRecord(ix->ev_end, b->stream)); ix->ev_end_set = true; }
        b->run_launches = launches; b->run_pending = true;
        return PGI_OK;
    } PGI_CATCH
}
static int batch_finish(pgi_batch *b) {
    if (!b) return fail(PGI_EINVAL, "null argument");
    if (!b->run_pending) return fail(PGI_EINVAL, "no step enqueued on this batch");
    try {
        DeviceGuard g(b->ix->device);
        const bool with_tags = b->ix->has_tags;
        int reruns = 0;
        while (wait_and_check(b, with_tags, !b->run_count_mode)) {
            if (++reruns > 8) { b->run_pending = false; return fail(PGI_ECUDA, "internal: the work buffers of a step did not converge"); }
            const int rc = batch_enqueue(b, b->run_count_mode, b->run_min_len, b->run_min_occ, b->run_mutex_held);
            if (rc != PGI_OK) { b->run_pending = false; return rc; }
        }
        b->run_pending = false;
        b->compact_done = b->want_compact;
        // [0] the K1 launch alone (device time between two in-stream events)
        b->stage_ms[0] = ev_ms(b, EV_K1A, EV_K1B); b->stage_ms[1] = ev_ms(b, EV_MEM, EV_ORDER); b->stage_ms[2] = ev_ms(b, EV_ORDER, EV_PLAN);
        b->stage_ms[3] = ev_ms(b, EV_PLAN, EV_DECODE); b->stage_ms[4] = ev_ms(b, EV_DECODE, EV_END); b->stage_ms[5] = ev_ms(b, EV_START, EV_END);
        b->stage_ms[6] = b->run_launches; b->stage_ms[7] = reruns;
        b->has_result = true;
        return PGI_OK;
    } catch (...) { b->run_pending = false; try { throw; } PGI_CATCH }
}
static int batch_run(pgi_batch *b, bool count_mode, uint64_t min_len, uint64_t min_occ, bool mutex_held = false) {
    const int rc = batch_enqueue(b, count_mode, min_len, min_occ, mutex_held);
    return rc == PGI_OK ? batch_finish(b) : rc;
}

extern "C" int pgi_batch_run_async(pgi_batch *b, uint64_t min_len, uint64_t min_occ) { return batch_enqueue(b, false, min_len, min_occ); }
extern "C" int pgi_batch_wait(pgi_batch *b) { return batch_finish(b); }
extern "C" int pgi_batch_run(pgi_batch *b, uint64_t min_len, uint64_t min_occ) { return batch_run(b, false, min_len, min_occ); }
extern "C" int pgi_batch_run_query_tags(pgi_batch *b) { return batch_run(b, true, 0, 0); }

extern "C" int pgi_batch_stage_ms(pgi_batch *b, double ms[8]) {
    if (!b || !ms) return fail(PGI_EINVAL, "null argument");
    for (int i = 0; i < 8; ++i) ms[i] = b->stage_ms[i];
    return PGI_OK;
}

extern "C" int pgi_batch_device_result(pgi_batch *b, pgi_result *out) {
    if (!b || !out) return fail(PGI_EINVAL, "null argument");
    if (!b->has_result) return fail(PGI_EINVAL, "no completed run on this batch");
    std::memset(out, 0, sizeof(*out));
    out->n_reads = b->n_reads; out->n_mems = b->n_mems; out->n_tags = b->n_tags; out->extensions = b->extensions; out->seed_lookups = b->seed_lookups; out->blocks_fetched = b->blocks_fetched;
    out->mem_off = b->d_mem_off.as<uint64_t>(); out->mems = b->d_mems.as<pgi_mem>(); out->tag_off = b->d_tag_off.as<uint64_t>();
    out->tags = b->d_tags.as<uint64_t>(); out->n_runs = b->d_n_runs.as<uint64_t>();
    return PGI_OK;
}

extern "C" int pgi_batch_fetch(pgi_batch *b, pgi_result *out) {
    if (!b || !out) return fail(PGI_EINVAL, "null argument");
    if (!b->has_result) return fail(PGI_EINVAL, "no completed run on this batch");
    try {
        DeviceGuard g(b->ix->device);
        const uint64_t n = b->n_reads, nm = b->n_mems, nt = b->n_tags; cudaStream_t st = b->stream;
        b->h_mem_off.need((n + 1) * 8); b->h_mems.need((nm + 1) * 32); b->h_tag_off.need((nm + 1) * 8);
        b->h_tags.need((nt + 1) * 8); b->h_n_runs.need((nm + 1) * 8);
        CK(cudaMemcpyAsync(b->h_mem_off.p, b->d_mem_off.p, (n + 1) * 8, cudaMemcpyDeviceToHost, st));
        if (nm) CK(cudaMemcpyAsync(b->h_mems.p, b->d_mems.p, nm * 32, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(b->h_tag_off.p, b->d_tag_off.p, (nm + 1) * 8, cudaMemcpyDeviceToHost, st));
        if (nt) CK(cudaMemcpyAsync(b->h_tags.p, b->d_tags.p, nt * 8, cudaMemcpyDeviceToHost, st));
        if (nm) CK(cudaMemcpyAsync(b->h_n_runs.p, b->d_n_runs.p, nm * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        std::memset(out, 0, sizeof(*out));
        out->n_reads = n; out->n_mems = nm; out->n_tags = nt; out->extensions = b->extensions; out->seed_lookups = b->seed_lookups; out->blocks_fetched = b->blocks_fetched;
        out->mem_off = b->h_mem_off.as<uint64_t>(); out->mems = b->h_mems.as<pgi_mem>(); out->tag_off = b->h_tag_off.as<uint64_t>();
        out->tags = b->h_tags.as<uint64_t>(); out->n_runs = b->h_n_runs.as<uint64_t>();
        return PGI_OK;
    } PGI_CATCH
}

// ------------------------------------------------------------------------------------------ compact wire form (kernels: above, before collect_and_check)

extern "C" int pgi_batch_upload_fixed(pgi_batch *b, const char *reads, uint64_t n_reads, uint64_t read_len) {
    if (!b || (!reads && n_reads)) return fail(PGI_EINVAL, "null argument");
    try {
        DeviceGuard g(b->ix->device);
        const uint64_t total = n_reads * read_len;
        b->d_reads.need(total + 16); b->d_read_off.need((n_reads + 1) * 8);
        if (total) CK(cudaMemcpyAsync(b->d_reads.p, reads, total, cudaMemcpyHostToDevice, b->stream));
        CK(cudaMemsetAsync((char *) b->d_reads.p + total, 0, 16, b->stream));
        iota_mul_kernel<<<grid_for(n_reads + 1, 256), 256, 0, b->stream>>>(n_reads + 1, read_len, b->d_read_off.as<uint64_t>());
        CK(cudaGetLastError());
        b->reads64 = b->d_reads.as<uint64_t>(); b->read_off = b->d_read_off.as<uint64_t>(); b->n_reads = n_reads; b->has_result = false; b->compact_done = false; b->reads_bytes = total;
        return PGI_OK;
    } PGI_CATCH
}

extern "C" int pgi_batch_fetch_compact(pgi_batch *b, pgi_result_compact *out) {
    if (!b || !out) return fail(PGI_EINVAL, "null argument");
    if (!b->has_result) return fail(PGI_EINVAL, "no completed run on this batch");
    try {
        DeviceGuard g(b->ix->device);
        const uint64_t n = b->n_reads, nm = b->n_mems, nt = b->n_tags; cudaStream_t st = b->stream;
        const unsigned tag_bytes = compact_tag_bytes(b->ix);
        b->h_c_cnt.need((n + 1) * 4); b->h_c_rec.need((nm + 1) * 16); b->h_c_rt.need((nm + 1) * 8); b->h_c_tags.need((nt + 1) * tag_bytes);
        if (!b->compact_done) {                             // a step run through pgi_batch_run: pack now (pgi_find_mems_compact packs inside the step)
            b->d_c_cnt.need((n + 1) * 4); b->d_c_rec.need((nm + 1) * 16); b->d_c_rt.need((nm + 1) * 8); b->d_c_tags.need((nt + 1) * tag_bytes + 32);
            unsigned long long *flags = b->d_ctr.as<unsigned long long>() + ST_NOFIT;
            CK(cudaMemsetAsync(flags, 0, 8, st));
            if (n) compact_reads_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, b->d_mem_off.as<uint64_t>(), b->d_c_cnt.as<uint32_t>(), flags);
            if (nm) compact_mems_kernel<<<grid_for(nm, 256), 256, 0, st>>>(nm, nullptr, b->d_mems.as<uint4>(), b->d_tag_off.as<uint64_t>(), b->d_n_runs.as<uint64_t>(),
                                                                         b->d_c_rec.as<uint64_t>(), b->d_c_rt.as<uint32_t>(), flags);
            if (nt) compact_tags_kernel<<<grid_for(nt / 4 + 1, 256), 256, 0, st>>>(nt / 4 + 1, nullptr, nm, b->d_tag_off.as<uint64_t>(), b->d_tags.as<uint64_t>(), tag_bytes, b->d_c_tags.as<uint32_t>(), flags);
            CK(cudaGetLastError());
            peek_kernel<<<1, 32, 0, st>>>((const uint64_t *) flags, b->h_ctr.as<uint64_t>() + 12, 1);
        }
        if (n) CK(cudaMemcpyAsync(b->h_c_cnt.p, b->d_c_cnt.p, n * 4, cudaMemcpyDeviceToHost, st));
        if (nm) { CK(cudaMemcpyAsync(b->h_c_rec.p, b->d_c_rec.p, nm * 16, cudaMemcpyDeviceToHost, st)); CK(cudaMemcpyAsync(b->h_c_rt.p, b->d_c_rt.p, nm * 8, cudaMemcpyDeviceToHost, st)); }
        if (nt) CK(cudaMemcpyAsync(b->h_c_tags.p, b->d_c_tags.p, nt * tag_bytes, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (b->h_ctr.as<volatile uint64_t>()[12]) return fail(PGI_EINVAL, "a value of this result does not fit the compact form (read of 2^24 bases or more, position or count too large): use pgi_find_mems");
        std::memset(out, 0, sizeof(*out));
        out->n_reads = n; out->n_mems = nm; out->n_tags = nt; out->tag_bytes = tag_bytes;
        out->mem_cnt = b->h_c_cnt.as<uint32_t>(); out->rec = b->h_c_rec.as<uint64_t>(); out->runs_tags = b->h_c_rt.as<uint32_t>(); out->tags = b->h_c_tags.as<uint8_t>();
        out->extensions = b->extensions; out->seed_lookups = b->seed_lookups; out->blocks_fetched = b->blocks_fetched;
        return PGI_OK;
    } PGI_CATCH
}

extern "C" int pgi_result_expand(const pgi_result_compact *c, pgi_result *out) {
    if (!c || !out) return fail(PGI_EINVAL, "null argument");
    std::memset(out, 0, sizeof(*out));
    const uint64_t n = c->n_reads, nm = c->n_mems, nt = c->n_tags;
    out->mem_off = (uint64_t *) std::malloc((n + 1) * 8); out->mems = (pgi_mem *) std::malloc((nm + 1) * sizeof(pgi_mem));
    out->tag_off = (uint64_t *) std::malloc((nm + 1) * 8); out->tags = (uint64_t *) std::malloc((nt + 1) * 8); out->n_runs = (uint64_t *) std::malloc((nm + 1) * 8);
    if (!out->mem_off || !out->mems || !out->tag_off || !out->tags || !out->n_runs) { pgi_result_expanded_free(out); return fail(PGI_ENOMEM, "out of host memory"); }
    uint64_t acc = 0;
    for (uint64_t i = 0; i < n; ++i) { out->mem_off[i] = acc; acc += c->mem_cnt[i]; }
    out->mem_off[n] = acc;
    uint64_t t = 0;
    for (uint64_t m = 0; m < nm; ++m) {
        const uint64_t w0 = c->rec[2 * m], w1 = c->rec[2 * m + 1];
        pgi_mem &x = out->mems[m];
        x.start = w0 >> 40; x.end = x.start + (w1 >> 40); x.bwt_start = w0 & 0xFFFFFFFFFFull; x.size = (int64_t) (w1 & 0xFFFFFFFFFFull);
        out->n_runs[m] = c->runs_tags[2 * m]; out->tag_off[m] = t; t += c->runs_tags[2 * m + 1];
    }
    out->tag_off[nm] = t;
    for (uint64_t i = 0; i < nt; ++i) { uint64_t v = 0; std::memcpy(&v, c->tags + i * c->tag_bytes, c->tag_bytes); out->tags[i] = v; }
    out->n_reads = n; out->n_mems = nm; out->n_tags = nt; out->extensions = c->extensions; out->seed_lookups = c->seed_lookups; out->blocks_fetched = c->blocks_fetched;
    return PGI_OK;
}
extern "C" void pgi_result_expanded_free(pgi_result *r) {
    if (!r) return;
    std::free(r->mem_off); std::free(r->mems); std::free(r->tag_off); std::free(r->tags); std::free(r->n_runs);
    std::memset(r, 0, sizeof(*r));
}

// ------------------------------------------------------------------------------------------ one-call API
static int run_shared(pgi_index *ix, pgi_batch *b, bool count_mode, uint64_t min_len, uint64_t min_occ) {
    if (chain_mode() != 0) return batch_run(b, count_mode, min_len, min_occ);
    std::lock_guard<std::mutex> token(ix->compute_mu);     // PGI_CHAIN=0: one step at a time, ordered on the host
    return batch_run(b, count_mode, min_len, min_occ, true);
}
static pgi_batch *pool_take(pgi_index *ix) {
    { std::lock_guard<std::mutex> l(ix->pool_mu); if (!ix->pool.empty()) { pgi_batch *b = ix->pool.back(); ix->pool.pop_back(); return b; } }
    pgi_batch *b = nullptr;
    return pgi_batch_create(ix, nullptr, &b) == PGI_OK ? b : nullptr;
}
static void pool_give(pgi_index *ix, pgi_batch *b) { std::lock_guard<std::mutex> l(ix->pool_mu); ix->pool.push_back(b); }

static int one_call(pgi_index *ix, bool count_mode, const char *reads, const uint64_t *read_off, uint64_t n_reads,
                    uint64_t min_len, uint64_t min_occ, void *cuda_stream, pgi_result *out) {
    if (!ix || !read_off || !out || (!reads && n_reads)) return fail(PGI_EINVAL, "null argument");
    pgi_batch *b = pool_take(ix);
    if (!b) return PGI_ECUDA;
    cudaStream_t own = b->stream;
    if (cuda_stream) b->stream = (cudaStream_t) cuda_stream;
    static const bool trace = std::getenv("PGI_TRACE") != nullptr;            // per-call timeline on stderr (host clock, ms)
    static const auto t_origin = std::chrono::steady_clock::now();
    auto now_ms = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_origin).count(); };
    const double t0 = trace ? now_ms() : 0;
    // One upload at a time: concurrent callers that share the host-to-device link evenly ALL finish late (three 316 MB
    // uploads side by side: 17.6 ms until the first kernel can start; one after the other: 5.7 ms), and the same for the
    // result downloads. In steady state the pipeline is compute bound either way; this shortens its ramp-up and drain.
    int rc;
    {
        std::lock_guard<std::mutex> up(ix->upload_mu);
        rc = pgi_batch_upload(b, reads, read_off, n_reads);
        if (rc == PGI_OK && cudaStreamSynchronize(b->stream) != cudaSuccess) rc = fail(PGI_ECUDA, "host to device copy failed");
    }
    double t1 = 0;
    // Concurrent callers pipeline instead of interleaving: the host->device copy is waited for OUTSIDE the compute token,
    // the kernels of one call at a time own the GPU (a persistent grid fills every SM anyway; two of them interleaved only
    // delay each other's short follow-up kernels), and the device->host copy again runs outside the token. In steady state
    // call k+1 uploads and call k-1 downloads while call k computes.
    if (trace) t1 = now_ms();
    if (rc == PGI_OK) { b->want_compact = false; rc = run_shared(ix, b, count_mode, min_len, min_occ); }
    const double t2 = trace ? now_ms() : 0;
    if (rc == PGI_OK) { std::lock_guard<std::mutex> down(ix->download_mu); rc = pgi_batch_fetch(b, out); }
    if (trace) std::fprintf(stderr, "[pgi] ctx %p reads %llu: enter %.2f uploaded %.2f run_done %.2f (mem kernel %.2f, stages %.2f ms) fetched %.2f\n", (void *) b,
                            (unsigned long long) n_reads, t0, t1, t2, b->stage_ms[0], b->stage_ms[5], now_ms());
    b->stream = own;
    if (rc != PGI_OK) { pool_give(ix, b); return rc; }
    out->owner = b;                                      // the result lives in the batch's pinned buffers
    return PGI_OK;
}

extern "C" int pgi_find_mems(pgi_index *ix, const char *reads, const uint64_t *read_off, uint64_t n_reads,
                             uint64_t min_len, uint64_t min_occ, void *cuda_stream, pgi_result *out) {
    return one_call(ix, false, reads, read_off, n_reads, min_len, min_occ, cuda_stream, out);
}
extern "C" int pgi_query_tags(pgi_index *ix, const char *reads, const uint64_t *read_off, uint64_t n_reads,
                              void *cuda_stream, pgi_result *out) {
    return one_call(ix, true, reads, read_off, n_reads, 0, 0, cuda_stream, out);
}

extern "C" int pgi_find_mems_compact(pgi_index *ix, const char *reads, const uint64_t *read_off, uint64_t read_len, uint64_t n_reads,
                                     uint64_t min_len, uint64_t min_occ, void *cuda_stream, pgi_result_compact *out) {
    if (!ix || !out || (!reads && n_reads)) return fail(PGI_EINVAL, "null argument");
    pgi_batch *b = pool_take(ix);
    if (!b) return PGI_ECUDA;
    cudaStream_t own = b->stream;
    if (cuda_stream) b->stream = (cudaStream_t) cuda_stream;
    int rc;
    {   // the same three-stage pipeline as pgi_find_mems: upload | compute token | download, one of each at a time
        std::lock_guard<std::mutex> up(ix->upload_mu);
        rc = read_off ? pgi_batch_upload(b, reads, read_off, n_reads) : pgi_batch_upload_fixed(b, reads, n_reads, read_len);
        if (rc == PGI_OK && cudaStreamSynchronize(b->stream) != cudaSuccess) rc = fail(PGI_ECUDA, "host to device copy failed");
    }
    if (rc == PGI_OK) { b->want_compact = true; rc = run_shared(ix, b, false, min_len, min_occ); }
    b->want_compact = false;
    if (rc == PGI_OK) { std::lock_guard<std::mutex> down(ix->download_mu); rc = pgi_batch_fetch_compact(b, out); }
    b->stream = own;
    if (rc != PGI_OK) { pool_give(ix, b); return rc; }
    out->owner = b;
    return PGI_OK;
}
extern "C" void pgi_result_compact_free(pgi_result_compact *r) {
    if (!r || !r->owner) return;
    pgi_batch *b = (pgi_batch *) r->owner;
    pool_give(b->ix, b);
    std::memset(r, 0, sizeof(*r));
}

extern "C" void pgi_result_free(pgi_result *r) {
    if (!r || !r->owner) return;
    pgi_batch *b = (pgi_batch *) r->owner;
    pool_give(b->ix, b);
    std::memset(r, 0, sizeof(*r));
}

extern "C" int pgi_tag_query_batch(pgi_index *ix, uint64_t n, const uint64_t *start, const uint64_t *end, pgi_result *out) {
    if (!ix || !out || ((!start || !end) && n)) return fail(PGI_EINVAL, "null argument");
    if (!ix->has_tags) return fail(PGI_EINVAL, "index was created without tags");
    pgi_batch *b = pool_take(ix);
    if (!b) return PGI_ECUDA;
    try {
        DeviceGuard g(ix->device);
        b->d_reads.need(std::max<uint64_t>(1, n) * 16);
        uint64_t *ds = b->d_reads.as<uint64_t>(), *de = ds + n;
        if (n) { CK(cudaMemcpyAsync(ds, start, n * 8, cudaMemcpyHostToDevice, b->stream)); CK(cudaMemcpyAsync(de, end, n * 8, cudaMemcpyHostToDevice, b->stream)); }
        for (int tries = 0;; ++tries) {
            if (tries > 8) { pool_give(ix, b); return fail(PGI_ECUDA, "internal: the work buffers of a tag query did not converge"); }
            int launches = 0;
            CK(cudaEventRecord(b->ev[EV_START], b->stream));
            CK(cudaMemsetAsync(b->d_ctr.p, 0, ST_WORDS * 8, b->stream));
            set_nm_kernel<<<1, 1, 0, b->stream>>>(b->d_ctr.as<unsigned long long>(), n);
            CK(cudaGetLastError());
            b->mems_cap = n; b->part_cap = 0; b->mem_slots = n;
            enqueue_tag_stage(b, n, nullptr, ds, de, launches);
            enqueue_collect(b, true);
            if (!wait_and_check(b, true, false)) break;
        }
        b->n_reads = 0; b->n_mems = n; b->extensions = 0; b->seed_lookups = 0; b->has_result = true;
        b->d_mem_off.need(8); CK(cudaMemsetAsync(b->d_mem_off.p, 0, 8, b->stream));
        b->d_mems.need(std::max<uint64_t>(1, n) * 32);
        b->read_off = nullptr;
        int rc = pgi_batch_fetch(b, out);
        if (rc != PGI_OK) { pool_give(ix, b); return rc; }
        out->mems = nullptr; out->mem_off = nullptr; out->owner = b;
        return PGI_OK;
    } catch (const CudaError &e) { pool_give(ix, b); return fail(PGI_ECUDA, std::string(e.what) + ": " + cudaGetErrorString(e.e)); }
}

// FastLocate::locate_encoded for a batch of BWT ranges.
extern "C" int pgi_locate_batch(pgi_index *ix, uint64_t n, const uint64_t *first, const uint64_t *second, pgi_locate_result *out) {
    if (!ix || !out || ((!first || !second) && n)) return fail(PGI_EINVAL, "null argument");
    if (!ix->has_ri) return fail(PGI_EINVAL, "index was created without an r-index (tags only)");
    if (!ix->has_locate) return fail(PGI_EFORMAT, ix->info.is_encoded && ix->info.sigma != 6
        ? "encoded r-index without N: the reference's locate_encoded skips six cumulatives where five exist (src/r-index.cpp:85-90,1313); not reproduced"
        : "the r-index file carries no suffix-array samples (samples / last / last_to_run are empty): nothing to locate with");
    std::memset(out, 0, sizeof(*out));
    std::vector<uint64_t> row_off(n + 1, 0);
    for (uint64_t i = 0; i < n; ++i) {
        if (second[i] >= first[i] && second[i] >= ix->info.n) return fail(PGI_EINVAL, "range beyond the BWT (undefined in the reference)");
        row_off[i + 1] = row_off[i] + (second[i] >= first[i] ? second[i] - first[i] + 1 : 0);
    }
    const uint64_t rows = row_off[n];
    pgi_batch *b = pool_take(ix);
    if (!b) return PGI_ECUDA;
    uint64_t *h_id_off = nullptr, *h_ids = nullptr, *h_row_off = nullptr, *h_sa = nullptr;
    try {
        DeviceGuard g(ix->device);
        cudaStream_t st = b->stream;
        DevBuf d_first, d_second, d_row_off, d_sa, d_ids, d_ids2, d_ucnt, d_id_off, d_out;
        struct Release { std::vector<DevBuf *> v; ~Release() { for (DevBuf *x : v) x->release(); } } rel{{&d_first, &d_second, &d_row_off, &d_sa, &d_ids, &d_ids2, &d_ucnt, &d_id_off, &d_out}};
        d_first.need((n + 1) * 8); d_second.need((n + 1) * 8); d_row_off.need((n + 1) * 8); d_sa.need((rows + 1) * 8); d_ids.need((rows + 1) * 8);
        d_ids2.need((rows + 1) * 8); d_ucnt.need((n + 1) * 8); d_id_off.need((n + 1) * 8);
        if (n) { CK(cudaMemcpyAsync(d_first.p, first, n * 8, cudaMemcpyHostToDevice, st)); CK(cudaMemcpyAsync(d_second.p, second, n * 8, cudaMemcpyHostToDevice, st)); }
        CK(cudaMemcpyAsync(d_row_off.p, row_off.data(), (n + 1) * 8, cudaMemcpyHostToDevice, st));
        CK(cudaMemsetAsync(d_ucnt.p, 0, (n + 1) * 8, st));
        uint64_t n_ids = 0;
        if (n && rows) {
            locate_kernel<<<grid_for(n, 128), 128, 0, st>>>(ix->loc, n, d_first.as<uint64_t>(), d_second.as<uint64_t>(), d_row_off.as<uint64_t>(), d_sa.as<uint64_t>(), d_ids.as<uint64_t>());
            CK(cudaGetLastError());
            size_t tmp = 0;
            CK(cub::DeviceSegmentedSort::SortKeys(nullptr, tmp, d_ids.as<uint64_t>(), d_ids2.as<uint64_t>(), (int64_t) rows, (int64_t) n,
                                                  d_row_off.as<int64_t>(), d_row_off.as<int64_t>() + 1, st));
            b->d_cub.need(tmp);
            CK(cub::DeviceSegmentedSort::SortKeys(b->d_cub.p, tmp, d_ids.as<uint64_t>(), d_ids2.as<uint64_t>(), (int64_t) rows, (int64_t) n,
                                                  d_row_off.as<int64_t>(), d_row_off.as<int64_t>() + 1, st));
            locate_unique_kernel<<<grid_for(n, 128), 128, 0, st>>>(n, d_row_off.as<uint64_t>(), d_ids2.as<uint64_t>(), d_ucnt.as<uint64_t>());
            CK(cudaGetLastError());
        }
        int launches = 0;
        exclusive_scan(b, d_ucnt.as<uint64_t>(), d_id_off.as<uint64_t>(), n + 1, launches);
        n_ids = read_back(b, d_id_off.as<uint64_t>() + n);
        d_out.need((n_ids + 1) * 8);
        if (n && n_ids) {
            locate_copy_kernel<<<grid_for(n, 128), 128, 0, st>>>(n, d_row_off.as<uint64_t>(), d_id_off.as<uint64_t>(), d_ids2.as<uint64_t>(), d_out.as<uint64_t>());
            CK(cudaGetLastError());
        }
        h_id_off = (uint64_t *) std::malloc((n + 1) * 8); h_ids = (uint64_t *) std::malloc((n_ids + 1) * 8);
        h_row_off = (uint64_t *) std::malloc((n + 1) * 8); h_sa = (uint64_t *) std::malloc((rows + 1) * 8);
        if (!h_id_off || !h_ids || !h_row_off || !h_sa) throw std::bad_alloc();
        CK(cudaMemcpyAsync(h_id_off, d_id_off.p, (n + 1) * 8, cudaMemcpyDeviceToHost, st));
        if (n_ids) CK(cudaMemcpyAsync(h_ids, d_out.p, n_ids * 8, cudaMemcpyDeviceToHost, st));
        if (rows) CK(cudaMemcpyAsync(h_sa, d_sa.p, rows * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        std::memcpy(h_row_off, row_off.data(), (n + 1) * 8);
        pool_give(ix, b);
        out->n = n; out->n_ids = n_ids; out->n_rows = rows; out->id_off = h_id_off; out->ids = h_ids; out->row_off = h_row_off; out->sa = h_sa;
        return PGI_OK;
    } catch (const CudaError &e) { pool_give(ix, b); std::free(h_id_off); std::free(h_ids); std::free(h_row_off); std::free(h_sa); return fail(PGI_ECUDA, std::string(e.what) + ": " + cudaGetErrorString(e.e));
    } catch (const std::bad_alloc &) { pool_give(ix, b); std::free(h_id_off); std::free(h_ids); std::free(h_row_off); std::free(h_sa); return fail(PGI_ENOMEM, "out of host memory"); }
}
extern "C" void pgi_locate_result_free(pgi_locate_result *r) {
    if (!r) return;
    std::free(r->id_off); std::free(r->ids); std::free(r->row_off); std::free(r->sa);
    std::memset(r, 0, sizeof(*r));
}

extern "C" int pgi_extend_batch(pgi_index *ix, uint64_t n, const pgi_bi_interval *in, const uint8_t *sym, const uint8_t *dir, pgi_bi_interval *out) {
    if (!ix || ((!in || !sym || !dir || !out) && n)) return fail(PGI_EINVAL, "null argument");
    if (!ix->has_ri) return fail(PGI_EINVAL, "index was created without an r-index (tags only)");
    if (!n) return PGI_OK;
    try {
        DeviceGuard g(ix->device);
        DevBuf d_in, d_sym, d_dir, d_out;
        d_in.need(n * 24); d_sym.need(n); d_dir.need(n); d_out.need(n * 24);
        CK(cudaMemcpy(d_in.p, in, n * 24, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_sym.p, sym, n, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_dir.p, dir, n, cudaMemcpyHostToDevice));
        fmd_extend_kernel<<<grid_for(n, 128), 128>>>(ix->ri, ix->d_desc, n, d_in.as<uint64_t>(), d_sym.as<uint8_t>(), d_dir.as<uint8_t>(), d_out.as<uint64_t>());
        CK(cudaGetLastError());
        CK(cudaMemcpy(out, d_out.p, n * 24, cudaMemcpyDeviceToHost));
        d_in.release(); d_sym.release(); d_dir.release(); d_out.release();
        return PGI_OK;
    } PGI_CATCH
}

extern "C" int pgi_rank_batch(pgi_index *ix, uint64_t n, const uint64_t *pos, uint64_t *out8) {
    if (!ix || ((!pos || !out8) && n)) return fail(PGI_EINVAL, "null argument");
    if (!ix->has_ri) return fail(PGI_EINVAL, "index was created without an r-index (tags only)");
    if (!n) return PGI_OK;
    try {
        DeviceGuard g(ix->device);
        DevBuf d_pos, d_out; d_pos.need(n * 8); d_out.need(n * 64);
        CK(cudaMemcpy(d_pos.p, pos, n * 8, cudaMemcpyHostToDevice));
        SlotTable st; for (int i = 0; i < 8; ++i) st.s[i] = ix->consts.slot[i]; st.sigma = (uint32_t) ix->consts.sigma;
        rank_kernel<<<grid_for(n, 128), 128>>>(ix->ri, st, n, d_pos.as<uint64_t>(), d_out.as<uint64_t>());
        CK(cudaGetLastError());
        CK(cudaMemcpy(out8, d_out.p, n * 64, cudaMemcpyDeviceToHost));
        d_pos.release(); d_out.release();
        return PGI_OK;
    } PGI_CATCH
}

// ------------------------------------------------------------------------------------------ allocation-free rank batch
// pgi_rank_batch for callers that may not allocate device memory (code under CUDA graph capture, serving loops, harnesses
// that forbid cudaMalloc). The two staging buffers pgi_rank_batch takes from cudaMalloc (positions in, rank vectors out)
// live in ONE file-scope __device__ array instead, which every GPU that loads libpgi.so holds until the process exits:
//   bytes [0, 8 * MAX)          positions
//   bytes [8 * MAX, 72 * MAX)   rank vectors, 8 u64 per position
// In host code `g_rank_scratch` is only the host-side stand-in of the symbol: its device address is taken once per device,
// with that device current, by cudaGetSymbolAddress, and that pointer is what the kernel receives. Every call shares the
// array (it keeps the previous call's contents; rank_kernel writes all 8 words of every position below n and nothing is
// read beyond them, so nothing has to start at zero, but the call clears the part of the output it is about to use anyway so
// that a kernel that did not run cannot pass for one that did). Calls are serialised inside one process by a mutex; calls
// from different processes do not share the array (one copy per context).
static __device__ __align__(256) unsigned char g_rank_scratch[(size_t) PGI_RANK_NOALLOC_MAX * 72];
static_assert(((size_t) PGI_RANK_NOALLOC_MAX * 8) % 256 == 0, "the output half of the scratch array stays 256-byte aligned");

extern "C" int pgi_rank_batch_noalloc(pgi_index *ix, uint64_t n, const uint64_t *pos, uint64_t *out8) {
    static std::mutex mu;
    static unsigned char *scratch_of[64];                   // device address of g_rank_scratch, per device
    if (!ix || ((!pos || !out8) && n)) return fail(PGI_EINVAL, "null argument");
    if (!ix->has_ri) return fail(PGI_EINVAL, "index was created without an r-index (tags only)");
    if (n > (uint64_t) PGI_RANK_NOALLOC_MAX)
        return fail(PGI_EINVAL, "pgi_rank_batch_noalloc: n = " + std::to_string(n) + " exceeds PGI_RANK_NOALLOC_MAX = " +
                    std::to_string((uint64_t) PGI_RANK_NOALLOC_MAX) + " (static scratch); split the batch or call pgi_rank_batch");
    if (ix->device < 0 || ix->device >= 64) return fail(PGI_EINVAL, "pgi_rank_batch_noalloc: device index beyond 63");
    if (!n) return PGI_OK;
    std::lock_guard<std::mutex> lock(mu);
    int prev = -1; cudaGetDevice(&prev);
    cudaError_t e = cudaSetDevice(ix->device);
    const char *what = "cudaSetDevice";
    if (e == cudaSuccess && !scratch_of[ix->device]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_rank_scratch); what = "cudaGetSymbolAddress(g_rank_scratch)";
        if (e == cudaSuccess) scratch_of[ix->device] = (unsigned char *) p;
    }
    if (e == cudaSuccess) {
        unsigned char *base = scratch_of[ix->device];
        uint64_t *d_pos = (uint64_t *) base, *d_out = (uint64_t *) (base + (size_t) PGI_RANK_NOALLOC_MAX * 8);
        SlotTable st; for (int i = 0; i < 8; ++i) st.s[i] = ix->consts.slot[i]; st.sigma = (uint32_t) ix->consts.sigma;
        e = cudaMemcpy(d_pos, pos, n * 8, cudaMemcpyHostToDevice); what = "cudaMemcpy(positions)";
        if (e == cudaSuccess) { e = cudaMemset(d_out, 0xff, n * 64); what = "cudaMemset(rank vectors)"; }
        if (e == cudaSuccess) {
            rank_kernel<<<grid_for(n, 128), 128>>>(ix->ri, st, n, d_pos, d_out);
            e = cudaGetLastError(); what = "rank_kernel launch";
        }
        if (e == cudaSuccess) { e = cudaMemcpy(out8, d_out, n * 64, cudaMemcpyDeviceToHost); what = "cudaMemcpy(rank vectors)"; }
    }
    if (prev >= 0) cudaSetDevice(prev);
    if (e != cudaSuccess) return fail(PGI_ECUDA, std::string("pgi_rank_batch_noalloc: ") + what + ": " + cudaGetErrorString(e));
    return PGI_OK;
}

// ------------------------------------------------------------------------------------------ text + microbenchmark
static inline void put_u64(std::string &s, uint64_t v) {
    char buf[24]; int i = 24;
    do { buf[--i] = (char) ('0' + v % 10); v /= 10; } while (v);
    s.append(buf + i, 24 - i);
}

// stdout grammar of src/find_mems.cpp:115-139 and src/tag_arrays.cpp:885-889. Host work, and the part of
// the find_mems PROGRAM that bounds it (the GPU produces results 20 x faster than text can be written): the reads of a result
// are cut into contiguous ranges, one per host thread this process may run on (PGI_FORMAT_THREADS overrides); a first pass
// computes the exact length of every range's text (decimal lengths only), the buffer is allocated once, and a second pass
// writes every range in place with raw pointer stores (two digits per division). Round 1 appended to one std::string on one
// thread (3 M reads/s); per-thread strings + a parallel copy managed 9.5 M reads/s on 16 threads.
static inline unsigned dec_len(uint64_t v) {
    static const uint64_t p10[20] = {1ull, 10ull, 100ull, 1000ull, 10000ull, 100000ull, 1000000ull, 10000000ull, 100000000ull, 1000000000ull, 10000000000ull,
                                     100000000000ull, 1000000000000ull, 10000000000000ull, 100000000000000ull, 1000000000000000ull, 10000000000000000ull,
                                     100000000000000000ull, 1000000000000000000ull, 10000000000000000000ull};
    if (v < 10) return 1;
    const unsigned t = ((64u - (unsigned) __builtin_clzll(v)) * 1233u) >> 12;      // floor(log10(v)) or one less
    return t + (v >= p10[t] ? 1u : 0u);
}
static inline unsigned dec_len_i(int64_t v) { return v < 0 ? 1u + dec_len((uint64_t) 0 - (uint64_t) v) : dec_len((uint64_t) v); }
static const char kDigits2[201] = "00010203040506070809101112131415161718192021222324252627282930313233343536373839404142434445464748495051525354555657585960616263646566676869707172737475767778798081828384858687888990919293949596979899";
static inline char *put_dec(char *p, uint64_t v) {
    const unsigned n = dec_len(v); char *e = p + n;
    while (v >= 100) { const unsigned q = (unsigned) (v % 100); v /= 100; e -= 2; std::memcpy(e, kDigits2 + 2 * q, 2); }
    if (v >= 10) { e -= 2; std::memcpy(e, kDigits2 + 2 * v, 2); } else *--e = (char) ('0' + v);
    return p + n;
}
static inline char *put_dec_i(char *p, int64_t v) { if (v < 0) { *p++ = '-'; return put_dec(p, (uint64_t) 0 - (uint64_t) v); } return put_dec(p, (uint64_t) v); }
template <size_t N> static inline char *put_lit(char *p, const char (&lit)[N]) { std::memcpy(p, lit, N - 1); return p + (N - 1); }

static size_t text_size_range(const pgi_result *r, uint64_t first_seq, int with_tags, uint64_t a, uint64_t b) {
    size_t n = 0;
    for (uint64_t i = a; i < b; ++i) {
        n += 5 + dec_len(first_seq + i) + 1 + 1;                                         // "Seq: " v "\n" ... "\n"
        for (uint64_t m = r->mem_off[i]; m < r->mem_off[i + 1]; ++m) {
            const pgi_mem &x = r->mems[m];
            n += 11 + dec_len(x.start) + 11 + dec_len(x.end) + 12 + dec_len(x.bwt_start) + 7 + dec_len_i(x.size) + 1;
            if (with_tags) {
                const uint64_t ta = r->tag_off[m], te = r->tag_off[m + 1];
                n += 28 + dec_len(te - ta) + 1 + 1;
                for (uint64_t t = ta; t < te; ++t) n += dec_len(r->tags[t]) + 2;
            }
        }
    }
    return n;
}
static char *format_range(const pgi_result *r, uint64_t first_seq, int with_tags, uint64_t a, uint64_t b, char *p) {
    for (uint64_t i = a; i < b; ++i) {
        p = put_lit(p, "Seq: "); p = put_dec(p, first_seq + i); *p++ = '\n';
        for (uint64_t m = r->mem_off[i]; m < r->mem_off[i + 1]; ++m) {
            const pgi_mem &x = r->mems[m];
            p = put_lit(p, "MEM START: "); p = put_dec(p, x.start); p = put_lit(p, ", MEM END: "); p = put_dec(p, x.end);
            p = put_lit(p, " BWT START: "); p = put_dec(p, x.bwt_start); p = put_lit(p, " SIZE: "); p = put_dec_i(p, x.size); *p++ = '\n';
            if (with_tags) {
                const uint64_t ta = r->tag_off[m], te = r->tag_off[m + 1];
                p = put_lit(p, "Number of unique positions: "); p = put_dec(p, te - ta); *p++ = '\n';
                for (uint64_t t = ta; t < te; ++t) { p = put_dec(p, r->tags[t]); *p++ = ','; *p++ = ' '; }
                *p++ = '\n';
            }
        }
        *p++ = '\n';
    }
    return p;
}
extern "C" char *pgi_format_text(const pgi_result *r, uint64_t first_seq, int with_tags, uint64_t *len) {
    if (!r || !len) { fail(PGI_EINVAL, "null argument"); return nullptr; }
    unsigned nt = 1;
    { cpu_set_t set; CPU_ZERO(&set); if (sched_getaffinity(0, sizeof(set), &set) == 0) nt = (unsigned) std::max(1, CPU_COUNT(&set)); }
    if (const char *e = std::getenv("PGI_FORMAT_THREADS")) nt = (unsigned) std::max(1, std::atoi(e));
    nt = (unsigned) std::min<uint64_t>(nt, std::max<uint64_t>(1, r->n_reads / 4096));
    std::vector<size_t> sz(nt), at(nt + 1);
    auto lo = [&](unsigned k) { return r->n_reads * k / nt; };
    auto each = [&](auto &&fn) {
        if (nt == 1) { fn(0u); return; }
        std::vector<std::thread> th;
        for (unsigned k = 0; k < nt; ++k) th.emplace_back([&fn, k] { fn(k); });
        for (auto &t : th) t.join();
    };
    char *buf = nullptr;
    try {
        each([&](unsigned k) { sz[k] = text_size_range(r, first_seq, with_tags, lo(k), lo(k + 1)); });
        at[0] = 0; for (unsigned k = 0; k < nt; ++k) at[k + 1] = at[k] + sz[k];
        buf = (char *) std::malloc(at[nt] + 1);
        if (!buf) { fail(PGI_ENOMEM, "out of host memory"); return nullptr; }
        each([&](unsigned k) { format_range(r, first_seq, with_tags, lo(k), lo(k + 1), buf + at[k]); });
    } catch (const std::exception &) { std::free(buf); fail(PGI_ENOMEM, "out of host memory"); return nullptr; }
    buf[at[nt]] = 0; *len = at[nt];
    return buf;
}
// stdout of src/query_tags.cpp:87-109 for a pgi_query_tags result: for every read with a match the two lines that
// query_compressed_compact prints itself (src/tag_arrays.cpp:885-889), then the summary line (:103-108).
extern "C" char *pgi_format_query_tags_text(const pgi_result *r, uint64_t first_index, uint64_t *len) {
    if (!r || !len) { fail(PGI_EINVAL, "null argument"); return nullptr; }
    std::string s;
    s.reserve((size_t) (r->n_reads * 96 + r->n_tags * 12));
    for (uint64_t i = 0; i < r->n_reads; ++i) {
        const pgi_mem &x = r->mems[i];
        if (x.size == 0) continue;                       // "Read i has no matches" goes to stderr in the reference
        const uint64_t a = r->tag_off[i], e = r->tag_off[i + 1];
        s += "Number of unique positions: "; put_u64(s, e - a); s.push_back('\n');
        for (uint64_t t = a; t < e; ++t) { put_u64(s, r->tags[t]); s += ", "; }
        s.push_back('\n');
        s += "read_index="; put_u64(s, first_index + i); s += "\tlen="; put_u64(s, x.end);
        s += "\tbwt_start="; put_u64(s, x.bwt_start); s += "\tbwt_end="; put_u64(s, x.bwt_start + (uint64_t) x.size - 1);
        s += "\truns="; put_u64(s, r->n_runs[i]); s.push_back('\n');
    }
    char *buf = (char *) std::malloc(s.size() + 1);
    if (!buf) { fail(PGI_ENOMEM, "out of host memory"); return nullptr; }
    std::memcpy(buf, s.data(), s.size()); buf[s.size()] = 0; *len = s.size();
    return buf;
}
extern "C" void pgi_free(void *p) { std::free(p); }

extern "C" void *pgi_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); fail(PGI_ENOMEM, "cudaHostAlloc failed"); return nullptr; }
    return p;
}
extern "C" void pgi_host_free(void *p) { if (p) cudaFreeHost(p); }

// cudaLimitMaxL2FetchGranularity of `device`: bytes the L2 fetches from DRAM around a missing sector (32, 64 or 128). A run
// block is one aligned 64-byte unit; with a granularity of 128 every block fetch drags in the neighbouring block as well.
extern "C" int pgi_l2_fetch_granularity(int device, int set_bytes, int *current) {
    if (int rc = check_device(device)) return rc;
    try {
        DeviceGuard g(device);
        if (set_bytes) CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t) set_bytes));
        size_t v = 0; CK(cudaDeviceGetLimit(&v, cudaLimitMaxL2FetchGranularity));
        if (current) *current = (int) v;
        return PGI_OK;
    } PGI_CATCH
}

extern "C" int pgi_random_access_gbs(int device, uint64_t buf_bytes, uint32_t bytes_per_access, uint64_t n_access, int chain, double *gbs) {
    if (!gbs || chain < 0 || (bytes_per_access != 32 && bytes_per_access != 64 && bytes_per_access != 128)) return fail(PGI_EINVAL, "bad argument");
    if (int rc = check_device(device)) return rc;
    try {
        DeviceGuard g(device);
        DevBuf buf, sink; buf.need(buf_bytes); sink.need(8);
        CK(cudaMemset(buf.p, 1, buf_bytes)); CK(cudaMemset(sink.p, 0, 8));
        const uint64_t n_slots = buf_bytes / bytes_per_access;
        if (n_slots < 1 || (chain <= 1 && n_slots >> 32)) return fail(PGI_EINVAL, "buffer too small or too large for the microbenchmark");
        if (chain == 0) {                                 // cooperative form: teams of lanes, 8 independent slots in flight per lane
            const uint32_t team = bytes_per_access / 16, rounds = 16;
            const uint64_t n_teams = std::max<uint64_t>(1, n_access / (rounds * 8ull));
            const unsigned grid = grid_for(n_teams * team, 256);
            const uint64_t accesses = (uint64_t) grid * 256 / team * rounds * 8ull;
            cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
            float best = 1e30f;
            for (int rep = 0; rep < 4; ++rep) {
                CK(cudaEventRecord(e0));
                unsigned long long *sk = sink.as<unsigned long long>();
                if (bytes_per_access == 32) gather_team_kernel<32><<<grid, 256>>>(buf.as<uint4>(), (uint32_t) n_slots, rounds, sk);
                else if (bytes_per_access == 64) gather_team_kernel<64><<<grid, 256>>>(buf.as<uint4>(), (uint32_t) n_slots, rounds, sk);
                else gather_team_kernel<128><<<grid, 256>>>(buf.as<uint4>(), (uint32_t) n_slots, rounds, sk);
                CK(cudaGetLastError());
                CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
                float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
                if (rep > 0 && ms < best) best = ms;
            }
            cudaEventDestroy(e0); cudaEventDestroy(e1);
            *gbs = (double) accesses * bytes_per_access / (best * 1e-3) / 1e9; buf.release(); sink.release();
            return PGI_OK;
        }
        uint64_t n_threads = std::max<uint64_t>(1, n_access / (uint64_t) chain);
        const uint32_t per_thread = 64;                   // independent form: 64 loads per thread
        if (chain == 1) n_threads = std::max<uint64_t>(256, (n_access / per_thread + 255) / 256 * 256);
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {              // rep 0 is the warm-up
            CK(cudaEventRecord(e0));
            const unsigned grid = grid_for(n_threads, 256);
            unsigned long long *sk = sink.as<unsigned long long>();
            if (chain == 1) {
                if (bytes_per_access == 32) gather_indep_kernel<32><<<grid, 256>>>(buf.as<uint4>(), (uint32_t) n_slots, per_thread, sk);
                else if (bytes_per_access == 64) gather_indep_kernel<64><<<grid, 256>>>(buf.as<uint4>(), (uint32_t) n_slots, per_thread, sk);
                else gather_indep_kernel<128><<<grid, 256>>>(buf.as<uint4>(), (uint32_t) n_slots, per_thread, sk);
            } else if (bytes_per_access == 32) gather_kernel<32><<<grid, 256>>>(buf.as<uint4>(), n_slots, n_threads, chain, sk);
            else if (bytes_per_access == 64) gather_kernel<64><<<grid, 256>>>(buf.as<uint4>(), n_slots, n_threads, chain, sk);
            else gather_kernel<128><<<grid, 256>>>(buf.as<uint4>(), n_slots, n_threads, chain, sk);
            CK(cudaGetLastError());
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
            float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep > 0 && ms < best) best = ms;
        }
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        if (chain == 1) { *gbs = (double) n_threads * per_thread * bytes_per_access / (best * 1e-3) / 1e9; buf.release(); sink.release(); return PGI_OK; }
        *gbs = (double) n_threads * chain * bytes_per_access / (best * 1e-3) / 1e9;
        buf.release(); sink.release();
        return PGI_OK;
    } PGI_CATCH
}
