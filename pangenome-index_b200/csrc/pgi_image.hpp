// pgi_image.hpp — the cached FLATTENED index image: what the reference's memory-mapped and chunked index managers
// (src/find_mems_memory_mapped.cpp:36-45, src/find_mems_chunked.cpp:40-50: open the .ri / .tags files by mmap, whole or in
// chunks of chunk_size_mb, instead of deserialising them into RAM) become on a GPU. The SDSL files are flattened ONCE
// into the device layout of pgi_layout.h and written as one file of page-aligned sections; opening an image maps it
// (mmap, read-only) and streams every section to HBM through two pinned staging buffers of `chunk` bytes — no parsing,
// no host copy of the index, host memory bounded by 2 x chunk. Host-only code (no CUDA in this header).
//
// File layout (little-endian):
//   ImageHeader | section payloads, each starting at a multiple of 4096
// The header carries every scalar / table of the index (RiConsts, the SparseRankConsts of the three sparse-rank
// structures, packed-vector shapes, tag scalars), the section table and FNV-1a checksums of the header and of every
// section (checked on open when asked to: PGI_IMAGE_VERIFY=1 or verify != 0).
#pragma once
#include <cerrno>
#include <cstring>
#include <string>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include "pgi_flatten.hpp"

namespace pgi {

enum ImageSection {
    SEC_BLOCKS = 0, SEC_Q_START, SEC_Q_VAL, SEC_RS_DIR, SEC_RS_LOW, SEC_LAST_DIR, SEC_LAST_LOW, SEC_SAMPLES, SEC_LAST_TO_RUN,
    SEC_IV_DIR, SEC_IV_LOW, SEC_STARTS, SEC_WORDS, SEC_STREAM, SEC_COUNT
};
static const uint32_t kSectionElem[SEC_COUNT] = {8, 8, 8, 8, 1, 8, 1, 8, 8, 8, 1, 8, 8, 1};

struct SectionEntry { uint64_t offset, bytes, checksum; };

#define PGI_IMAGE_MAGIC 0x31474D4949475050ull   /* "PPGIIMG1" */
#define PGI_IMAGE_VERSION 3u   /* 2: crowded buckets as arithmetic sub-buckets, bucket shift up to 12; 3: block headers in limb form */
#define PGI_IMAGE_ALIGN 4096u

// Everything of an index that is not a big array. POD, written verbatim.
struct ImageMeta {
    RiConsts k;
    uint64_t n_file_runs, n_ref_blocks;
    uint32_t encoded, has_ri, has_tags, has_locate;
    SparseRankConsts rs_k, last_k, iv_k;
    uint64_t samples_n, samples_width, ltr_n, ltr_width;
    uint64_t tag_layout, tag_n_items, tag_width, tag_starts_regular;
};

struct ImageHeader {
    uint64_t magic; uint32_t version, header_bytes, meta_bytes, n_sections, block_u32, runs_per_block;
    uint64_t file_bytes;
    uint64_t src_ri_bytes, src_tag_bytes;       // sizes of the SDSL files the image was made from (staleness check of the CLIs)
    uint64_t src_ri_sum, src_tag_sum;           // FNV-1a of their first and last 64 KiB
    ImageMeta meta;
    SectionEntry sec[SEC_COUNT];
    uint64_t header_checksum;                   // FNV-1a of all bytes before this field
};

inline uint64_t fnv1a(const void *p, size_t n, uint64_t h = 0xcbf29ce484222325ull) {
    // 8 bytes per step (the byte-wise form is 1 GB/s; this one keeps image verification near memory speed)
    const uint8_t *b = (const uint8_t *) p; size_t i = 0;
    for (; i + 8 <= n; i += 8) { uint64_t w; std::memcpy(&w, b + i, 8); h = (h ^ w) * 0x100000001b3ull; h ^= h >> 29; }
    for (; i < n; ++i) h = (h ^ b[i]) * 0x100000001b3ull;
    return h;
}
// Content stamp of a source file (staleness check of the cached image): FNV-1a over the WHOLE file -- it runs near memory
// speed, and a stamp of the edges only would accept a file that was rebuilt with the same size and changed in the middle.
inline uint64_t edge_sum(const void *p, size_t n) { return fnv1a(p, n); }

// A view of a flattened index: scalars + one (pointer, bytes) pair per section. Built either over the vectors of a
// fresh RiImage / TagImage or over a mapped image file; the uploader (pgi_capi.cu) only sees this.
struct IndexView {
    ImageMeta m;
    const void *ptr[SEC_COUNT]; uint64_t bytes[SEC_COUNT];
};

inline void view_of(const RiImage *ri, const TagImage *tg, IndexView &v) {
    std::memset(&v, 0, sizeof(v));
    auto set = [&](int s, const void *p, uint64_t bytes) { v.ptr[s] = p; v.bytes[s] = bytes; };
    if (ri) {
        v.m.k = ri->k; v.m.n_file_runs = ri->n_file_runs; v.m.n_ref_blocks = ri->n_ref_blocks; v.m.encoded = ri->encoded; v.m.has_ri = 1;
        set(SEC_BLOCKS, ri->blocks.data(), ri->blocks.size() * 8);
        set(SEC_Q_START, ri->q_start.data(), ri->q_start.size() * 8); set(SEC_Q_VAL, ri->q_val.data(), ri->q_val.size() * 8);
        if (ri->k.locate_ok) {
            v.m.has_locate = 1; v.m.rs_k = ri->run_starts.k; v.m.last_k = ri->last.k;
            set(SEC_RS_DIR, ri->run_starts.dir.data(), ri->run_starts.dir.size() * 8); set(SEC_RS_LOW, ri->run_starts.low.data(), ri->run_starts.low.size());
            set(SEC_LAST_DIR, ri->last.dir.data(), ri->last.dir.size() * 8); set(SEC_LAST_LOW, ri->last.low.data(), ri->last.low.size());
            set(SEC_SAMPLES, ri->samples.words.data(), ri->samples.words.size() * 8); v.m.samples_n = ri->samples.n; v.m.samples_width = ri->samples.width;
            set(SEC_LAST_TO_RUN, ri->last_to_run.words.data(), ri->last_to_run.words.size() * 8); v.m.ltr_n = ri->last_to_run.n; v.m.ltr_width = ri->last_to_run.width;
        }
    }
    if (tg) {
        v.m.has_tags = 1; v.m.iv_k = tg->intervals.k; v.m.tag_layout = (uint64_t) tg->layout; v.m.tag_n_items = tg->n_items;
        v.m.tag_width = tg->width; v.m.tag_starts_regular = tg->starts_regular;
        set(SEC_IV_DIR, tg->intervals.dir.data(), tg->intervals.dir.size() * 8); set(SEC_IV_LOW, tg->intervals.low.data(), tg->intervals.low.size());
        set(SEC_STARTS, tg->starts.data(), tg->starts.size() * 8); set(SEC_WORDS, tg->words.data(), tg->words.size() * 8);
        set(SEC_STREAM, tg->stream.data(), tg->stream.size());
    }
}

inline void write_image(const IndexView &v, const char *path, uint64_t src_ri_bytes, uint64_t src_ri_sum, uint64_t src_tag_bytes, uint64_t src_tag_sum) {
    ImageHeader h; std::memset(&h, 0, sizeof(h));
    h.magic = PGI_IMAGE_MAGIC; h.version = PGI_IMAGE_VERSION; h.header_bytes = sizeof(ImageHeader); h.meta_bytes = sizeof(ImageMeta);
    h.n_sections = SEC_COUNT; h.block_u32 = PGI_BLOCK_U32; h.runs_per_block = PGI_RUNS_PER_BLOCK;
    h.src_ri_bytes = src_ri_bytes; h.src_ri_sum = src_ri_sum; h.src_tag_bytes = src_tag_bytes; h.src_tag_sum = src_tag_sum;
    h.meta = v.m;
    auto align = [](uint64_t x) { return (x + PGI_IMAGE_ALIGN - 1) / PGI_IMAGE_ALIGN * PGI_IMAGE_ALIGN; };
    uint64_t at = align(sizeof(ImageHeader));
    for (int s = 0; s < SEC_COUNT; ++s) {
        h.sec[s].offset = at; h.sec[s].bytes = v.bytes[s]; h.sec[s].checksum = fnv1a(v.ptr[s], v.bytes[s]);
        at = align(at + v.bytes[s]);
    }
    h.file_bytes = at;
    h.header_checksum = fnv1a(&h, offsetof(ImageHeader, header_checksum));
    const std::string tmp = std::string(path) + ".tmp";
    const int fd = ::open(tmp.c_str(), O_CREAT | O_TRUNC | O_WRONLY, 0644);
    if (fd < 0) throw std::runtime_error("cannot create " + tmp + ": " + std::strerror(errno));
    auto put = [&](const void *p, uint64_t n, uint64_t off) {
        const uint8_t *b = (const uint8_t *) p;
        while (n) {
            const ssize_t w = ::pwrite(fd, b, (size_t) std::min<uint64_t>(n, 1u << 30), (off_t) off);
            if (w <= 0) { const int e = errno; ::close(fd); ::unlink(tmp.c_str()); throw std::runtime_error("write to " + tmp + " failed: " + std::strerror(e)); }
            b += w; n -= (uint64_t) w; off += (uint64_t) w;
        }
    };
    put(&h, sizeof(h), 0);
    for (int s = 0; s < SEC_COUNT; ++s) put(v.ptr[s], v.bytes[s], h.sec[s].offset);
    if (::ftruncate(fd, (off_t) h.file_bytes) != 0 || ::close(fd) != 0 || ::rename(tmp.c_str(), path) != 0) {
        const int e = errno; ::unlink(tmp.c_str()); throw std::runtime_error(std::string("cannot finish ") + path + ": " + std::strerror(e));
    }
}

// Read-only mapping of an image file. The header is validated on open (magic, version, struct sizes, layout constants,
// section bounds, sizes implied by the scalars); payload checksums only on request (they read the whole file).
struct MappedImage {
    const uint8_t *base = nullptr; uint64_t len = 0; const ImageHeader *h = nullptr;
    MappedImage() = default;
    MappedImage(const MappedImage &) = delete; MappedImage &operator=(const MappedImage &) = delete;
    ~MappedImage() { if (base) ::munmap((void *) base, len); }

    void open(const char *path, bool verify_payload) {
        const int fd = ::open(path, O_RDONLY);
        if (fd < 0) throw FormatError(std::string("cannot open image ") + path + ": " + std::strerror(errno));
        struct stat st;
        if (::fstat(fd, &st) != 0 || (uint64_t) st.st_size < sizeof(ImageHeader)) { ::close(fd); throw FormatError("image file is too short"); }
        len = (uint64_t) st.st_size;
        void *p = ::mmap(nullptr, len, PROT_READ, MAP_SHARED, fd, 0);
        ::close(fd);
        if (p == MAP_FAILED) { len = 0; throw FormatError(std::string("mmap of the image failed: ") + std::strerror(errno)); }
        base = (const uint8_t *) p; h = (const ImageHeader *) base;
        if (h->magic != PGI_IMAGE_MAGIC) throw FormatError("not a pgi index image (bad magic)");
        if (h->version != PGI_IMAGE_VERSION || h->header_bytes != sizeof(ImageHeader) || h->meta_bytes != sizeof(ImageMeta) || h->n_sections != SEC_COUNT ||
            h->block_u32 != PGI_BLOCK_U32 || h->runs_per_block != PGI_RUNS_PER_BLOCK)
            throw FormatError("index image was written by an incompatible build (version / layout constants differ)");
        if (h->header_checksum != fnv1a(h, offsetof(ImageHeader, header_checksum))) throw FormatError("index image header is corrupt (checksum)");
        if (h->file_bytes != len) throw FormatError("index image is truncated");
        for (int s = 0; s < SEC_COUNT; ++s) {
            const SectionEntry &e = h->sec[s];
            if (e.offset % PGI_IMAGE_ALIGN || e.offset > len || e.bytes > len - e.offset || e.bytes % kSectionElem[s]) throw FormatError("index image section table is corrupt");
        }
        const ImageMeta &m = h->meta;
        if (m.has_ri && h->sec[SEC_BLOCKS].bytes != m.k.n_blocks * 64) throw FormatError("index image: block section does not match n_blocks");
        if (m.has_ri && (h->sec[SEC_Q_START].bytes != m.k.n_quirk * 8 || h->sec[SEC_Q_VAL].bytes != m.k.n_quirk * 8)) throw FormatError("index image: quirk tables do not match n_quirk");
        auto sr_ok = [&](const SparseRankConsts &k, int dir, int low) {
            return h->sec[dir].bytes == (k.n_buckets + 1) * 8 && h->sec[low].bytes >= k.m * k.low_bytes;
        };
        if (m.has_locate && (!sr_ok(m.rs_k, SEC_RS_DIR, SEC_RS_LOW) || !sr_ok(m.last_k, SEC_LAST_DIR, SEC_LAST_LOW) ||
                             h->sec[SEC_SAMPLES].bytes < (m.samples_n * m.samples_width + 7) / 8 || h->sec[SEC_LAST_TO_RUN].bytes < (m.ltr_n * m.ltr_width + 7) / 8))
            throw FormatError("index image: locate sections do not match their scalars");
        if (m.has_tags && !sr_ok(m.iv_k, SEC_IV_DIR, SEC_IV_LOW)) throw FormatError("index image: tag interval sections do not match their scalars");
        if (m.has_tags && m.tag_layout == 0 && h->sec[SEC_WORDS].bytes < (m.tag_n_items * m.tag_width + 7) / 8) throw FormatError("index image: tag items do not match their scalars");
        if (m.has_tags && m.tag_layout == 1 && h->sec[SEC_STREAM].bytes < m.tag_n_items) throw FormatError("index image: tag stream does not match its scalars");
        if (verify_payload)
            for (int s = 0; s < SEC_COUNT; ++s)
                if (fnv1a(base + h->sec[s].offset, h->sec[s].bytes) != h->sec[s].checksum) throw FormatError("index image payload is corrupt (section checksum)");
    }
    void view(IndexView &v) const {
        std::memset(&v, 0, sizeof(v)); v.m = h->meta;
        for (int s = 0; s < SEC_COUNT; ++s) { v.ptr[s] = base + h->sec[s].offset; v.bytes[s] = h->sec[s].bytes; }
    }
};

} // namespace pgi
