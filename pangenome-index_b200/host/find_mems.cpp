// find_mems — batched, multi-GPU driver with the reference's command line and stdout grammar.
//
//   find_mems <r_index.ri> <tags> <reads.txt> <min_mem_length> <min_occ> [--gpus N] [--batch READS]
//             [--tags-layout auto|compact|bytecode] [--no-text]
//
// Replaces src/find_mems.cpp:13-147 of the reference (one read at a time, one thread): reads are cut into batches of
// non-empty lines, batches are dealt to the GPUs (index replicated per GPU, two batches in flight per GPU so that the
// copies of one overlap the kernels of the other), every batch goes through ONE pgi_find_mems call (MEMs + tags), the
// text of src/find_mems.cpp:115-139 / src/tag_arrays.cpp:885-889 is produced by pgi_format_text and written in read
// order. Sequence numbers count non-empty lines from 1, like the reference's `i`. No CPU query path exists here.
//
// Compiled with -DPGI_QUERY_TAGS the same driver is `query_tags <r_index.ri> <tags> <reads.txt> [options]`, replacing
// src/query_tags.cpp:39-112: every batch goes through pgi_query_tags (whole-read count_encoded / count + the tag query
// of the final range) and the text is that of src/query_tags.cpp:103-108 preceded by the two lines
// query_compressed_compact prints itself; read_index counts non-empty lines from 0 (readSequencesFromFile :27-37).
//
// Compiled with -DPGI_MMAP / -DPGI_CHUNKED it is `find_mems_memory_mapped <ri> <tags> <reads> <mem_length> <min_occ>`
// (src/find_mems_memory_mapped.cpp:13-114) / `find_mems_chunked ... [chunk_size_mb]` (src/find_mems_chunked.cpp:13-121;
// default 2048). The reference's variants avoid deserialising the index into RAM: MemoryMappedIndexManager maps the two
// files, the chunked manager maps them in windows of chunk_size_mb. Here the SDSL files are flattened ONCE into a cached
// device image next to the .ri file (<ri>.pgimg, or --image PATH; rebuilt when the source files change), and every later
// run maps that image and streams it to HBM through two pinned staging buffers -- of chunk_size_mb for the chunked
// variant, 64 MiB for the memory-mapped one: no parsing, host memory bounded by 2 x chunk. Like the reference's variants
// (FastLocate::load + TagArray::load_compressed_tags + query_compressed) they expect ByteCode tags, but take either
// layout. The per-read text is that of find_mems; the trailer adds the reference's size / chunk lines.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "panindexer.hpp"
#if defined(PGI_MMAP) || defined(PGI_CHUNKED)
#define PGI_IMAGE_DRIVER 1
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#endif

namespace {

#ifdef PGI_IMAGE_DRIVER
struct MapFile {                                 // read-only mapping of a source file (only the pages that are used get read)
    const char *p = nullptr; size_t len = 0;
    bool open(const std::string &path) {
        const int fd = ::open(path.c_str(), O_RDONLY);
        if (fd < 0) return false;
        struct stat st;
        if (::fstat(fd, &st) != 0) { ::close(fd); return false; }
        len = (size_t) st.st_size;
        if (len) { void *m = ::mmap(nullptr, len, PROT_READ, MAP_SHARED, fd, 0); if (m == MAP_FAILED) { ::close(fd); return false; } p = (const char *) m; }
        ::close(fd); return true;
    }
    ~MapFile() { if (p) ::munmap((void *) p, len); }
};
#endif

struct Batch {
    uint64_t id = 0, first_seq = 1, n = 0, bytes = 0;
    char *reads = nullptr; uint64_t *off = nullptr; size_t reads_cap = 0, off_cap = 0;
    void reserve_reads(size_t want) {
        if (want <= reads_cap) return;
        size_t cap = want + want / 2 + 4096; char *p = (char *) pgi_host_alloc(cap);
        if (!p) { std::fprintf(stderr, "find_mems: out of pinned memory\n"); std::exit(EXIT_FAILURE); }
        if (reads) { std::memcpy(p, reads, bytes); pgi_host_free(reads); }
        reads = p; reads_cap = cap;
    }
    void reserve_off(size_t want) {
        if (want <= off_cap) return;
        size_t cap = want + want / 2 + 64; uint64_t *p = (uint64_t *) pgi_host_alloc(cap * 8);
        if (!p) { std::fprintf(stderr, "find_mems: out of pinned memory\n"); std::exit(EXIT_FAILURE); }
        if (off) { std::memcpy(p, off, (n + 1) * 8); pgi_host_free(off); }
        off = p; off_cap = cap;
    }
    void clear() { n = 0; bytes = 0; reserve_off(2); off[0] = 0; }
    uint64_t expect = 0;                           // reads this batch will hold: the first read sizes the pinned buffers once
    void add(const char *s, size_t len) {          // (growing them geometrically re-allocated pinned memory a dozen times per batch)
        if (!reads_cap && expect) { reserve_reads((size_t) expect * (len + 2) + 4096); reserve_off((size_t) expect + 2); }
        reserve_reads(bytes + len + 16); reserve_off(n + 2);
        std::memcpy(reads + bytes, s, len); bytes += len; off[++n] = bytes;
    }
};

template <typename T> class Channel {          // small blocking queue
public:
    void push(T v) { { std::lock_guard<std::mutex> l(mu_); q_.push_back(v); } cv_.notify_one(); }
    bool pop(T &out) {
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [&] { return !q_.empty() || closed_; });
        if (q_.empty()) return false;
        out = q_.front(); q_.erase(q_.begin()); return true;
    }
    void close() { { std::lock_guard<std::mutex> l(mu_); closed_ = true; } cv_.notify_all(); }
private:
    std::mutex mu_; std::condition_variable cv_; std::vector<T> q_; bool closed_ = false;
};

[[maybe_unused]] std::string slurp_file(const std::string &path, const char *what) {
    std::ifstream in(path, std::ios::binary);
    if (!in) { std::fprintf(stderr, "Cannot open %s: %s\n", what, path.c_str()); std::exit(EXIT_FAILURE); }
    return panindexer::slurp(in);
}

} // namespace

int main(int argc, char **argv) {
#ifdef PGI_QUERY_TAGS
    const int n_fixed = 4;
    if (argc < n_fixed) {
        std::fprintf(stderr, "Usage: %s <r_index.ri> <compressed_tags.tags> <reads.txt> [--gpus N] [--batch READS] [--tags-layout auto|compact|bytecode] [--no-text]\n", argv[0]);
        return 1;
    }
    const uint64_t mem_length = 0, min_occ = 0;
#else
    const int n_fixed = 6;
    if (argc < n_fixed) {
        std::fprintf(stderr, "usage: %s <r_index.ri> <tags> <reads.txt> <min_mem_length> <min_occ> [--gpus N] [--batch READS] [--tags-layout auto|compact|bytecode] [--no-text]\n", argv[0]);
        return EXIT_FAILURE;
    }
    const uint64_t mem_length = (uint64_t) std::stoi(argv[4]), min_occ = (uint64_t) std::stoi(argv[5]);   // src/find_mems.cpp:17-18
#endif
    const std::string r_index_file = argv[1], tag_array_index = argv[2], reads_file = argv[3];
    int gpus = 1; uint64_t batch_reads = 1u << 20; std::string layout = "auto"; bool text = true;
    int first_opt = n_fixed;
#ifdef PGI_IMAGE_DRIVER
    std::string image_path = r_index_file + ".pgimg"; bool rebuild = false, verify = false;
    uint64_t chunk_size_mb = 64;
#ifdef PGI_CHUNKED
    chunk_size_mb = 2048;                                                                                // src/find_mems_chunked.cpp:29
    if (argc > n_fixed && std::strncmp(argv[n_fixed], "--", 2) != 0) { chunk_size_mb = (uint64_t) std::stoi(argv[n_fixed]); first_opt = n_fixed + 1; }
    if (chunk_size_mb < 1) chunk_size_mb = 1;
#endif
#endif
    for (int i = first_opt; i < argc; ++i) {
        std::string a = argv[i];
#ifdef PGI_IMAGE_DRIVER
        if (a == "--image" && i + 1 < argc) { image_path = argv[++i]; continue; }
        if (a == "--rebuild-image") { rebuild = true; continue; }
        if (a == "--verify-image") { verify = true; continue; }
#endif
        if (a == "--gpus" && i + 1 < argc) gpus = std::atoi(argv[++i]);
        else if (a == "--batch" && i + 1 < argc) batch_reads = std::strtoull(argv[++i], nullptr, 10);
        else if (a == "--tags-layout" && i + 1 < argc) layout = argv[++i];
        else if (a == "--no-text") text = false;
        else { std::fprintf(stderr, "find_mems: unknown option %s\n", a.c_str()); return EXIT_FAILURE; }
    }
    const int have = pgi_device_count();
    if (have < 1) { std::fprintf(stderr, "find_mems: no CUDA device (this build has no CPU path)\n"); return EXIT_FAILURE; }
    if (gpus < 1 || gpus > have) { std::fprintf(stderr, "find_mems: --gpus %d but %d device(s) visible\n", gpus, have); return EXIT_FAILURE; }
    if (batch_reads < 1) batch_reads = 1;

    using clk = std::chrono::steady_clock;
    auto t_load = clk::now();
#ifdef PGI_IMAGE_DRIVER
#ifdef PGI_CHUNKED
    std::cerr << "Opening chunked memory-mapped indices..." << std::endl;
    std::cerr << "Chunk size: " << chunk_size_mb << " MB" << std::endl;
#else
    std::cerr << "Opening memory-mapped indices..." << std::endl;
#endif
    std::vector<pgi_index *> ix(gpus, nullptr);
    pgi_image_info_t img;
    {
        MapFile ri, tg;
        if (!ri.open(r_index_file) || !tg.open(tag_array_index)) { std::cerr << "Failed to open memory-mapped indices" << std::endl; return EXIT_FAILURE; }
        const uint64_t ri_sum = pgi_image_source_sum(ri.p, ri.len), tg_sum = pgi_image_source_sum(tg.p, tg.len);
        bool fresh = !rebuild && pgi_image_info(image_path.c_str(), 0, &img) == PGI_OK && img.src_ri_bytes == ri.len && img.src_tag_bytes == tg.len &&
                     img.src_ri_sum == ri_sum && img.src_tag_sum == tg_sum &&
                     (layout == "auto" || img.tag_layout == (layout == "bytecode" ? (uint64_t) PGI_TAGS_BYTECODE_LEN : (uint64_t) PGI_TAGS_SDSL_COMPACT));
        if (!fresh) {
            std::cerr << "Flattening the index into " << image_path << " (first use, or the source files changed)" << std::endl;
            int rc = PGI_EFORMAT;                  // the reference's variants read ByteCode tags (load_compressed_tags): try that layout first
            if (layout == "bytecode" || layout == "auto") rc = pgi_image_write(ri.p, ri.len, tg.p, tg.len, PGI_TAGS_BYTECODE_LEN, image_path.c_str());
            if (rc == PGI_EFORMAT && (layout == "compact" || layout == "auto")) rc = pgi_image_write(ri.p, ri.len, tg.p, tg.len, PGI_TAGS_SDSL_COMPACT, image_path.c_str());
            if (rc != PGI_OK) { std::fprintf(stderr, "find_mems: cannot load the index: %s\n", pgi_last_error()); return EXIT_FAILURE; }
        } else std::cerr << "Using the cached image " << image_path << std::endl;
    }
    if (pgi_index_open_image(image_path.c_str(), 0, chunk_size_mb << 20, verify ? 1 : 0, &ix[0], &img) != PGI_OK) {
        std::fprintf(stderr, "find_mems: cannot open the index image: %s\n", pgi_last_error()); return EXIT_FAILURE;
    }
    for (int g = 1; g < gpus; ++g)
        if (pgi_index_clone_to(ix[0], g, &ix[g]) != PGI_OK) { std::fprintf(stderr, "find_mems: cannot replicate the index on GPU %d: %s\n", g, pgi_last_error()); return EXIT_FAILURE; }
    const double image_gb = (double) img.file_bytes / (1024.0 * 1024.0 * 1024.0);
    std::cerr << "Opening memory-mapped indices took " << std::chrono::duration<double>(clk::now() - t_load).count() << " seconds ("
              << img.chunks_copied << " chunk copies)" << std::endl;
    std::cerr << "Total memory-mapped size: " << image_gb << " GB" << std::endl;
#else
    std::cerr << "Reading the rindex file (encoded)" << std::endl;
    std::string ri = slurp_file(r_index_file, "r-index");
    std::cerr << "Reading the tag array index" << std::endl;
    std::string tg = slurp_file(tag_array_index, "tag arrays");
    std::vector<pgi_index *> ix(gpus, nullptr);
    {
        int rc = PGI_EFORMAT;
        if (layout == "compact" || layout == "auto") rc = pgi_index_create(ri.data(), ri.size(), tg.data(), tg.size(), PGI_TAGS_SDSL_COMPACT, 0, &ix[0]);
        if (rc == PGI_EFORMAT && (layout == "bytecode" || layout == "auto")) rc = pgi_index_create(ri.data(), ri.size(), tg.data(), tg.size(), PGI_TAGS_BYTECODE_LEN, 0, &ix[0]);
        if (rc != PGI_OK) { std::fprintf(stderr, "find_mems: cannot load the index: %s\n", pgi_last_error()); return EXIT_FAILURE; }
        for (int g = 1; g < gpus; ++g)
            if (pgi_index_clone_to(ix[0], g, &ix[g]) != PGI_OK) { std::fprintf(stderr, "find_mems: cannot replicate the index on GPU %d: %s\n", g, pgi_last_error()); return EXIT_FAILURE; }
    }
    ri.clear(); ri.shrink_to_fit(); tg.clear(); tg.shrink_to_fit();
    std::cerr << "Loading and flattening the index took " << std::chrono::duration<double>(clk::now() - t_load).count() << " seconds" << std::endl;

#endif

    FILE *rf = std::fopen(reads_file.c_str(), "rb");
    if (!rf) { std::cerr << "Cannot open reads file: " << reads_file << std::endl; return EXIT_FAILURE; }

    const int n_workers = 2 * gpus, n_batches = n_workers + 2;
    std::vector<Batch> pool(n_batches);
    Channel<Batch *> free_q, work_q;
    for (auto &b : pool) free_q.push(&b);
    std::mutex out_mu; std::condition_variable out_cv; std::map<uint64_t, std::pair<char *, uint64_t>> done;   // batch id -> text
    std::atomic<bool> failed{false}; std::atomic<uint64_t> total_batches{~0ull};
    double gpu_seconds = 0; std::mutex stat_mu; uint64_t total_reads = 0, total_mems = 0, total_tags = 0;

    // reader: non-empty lines -> batches (std::getline semantics: '\n' separated, no other trimming)
    std::thread reader([&] {
        std::vector<char> buf(1 << 24); std::string carry; uint64_t id = 0, seq = 1; Batch *cur = nullptr;
        auto begin = [&]() { if (!free_q.pop(cur)) return false; cur->expect = batch_reads; cur->clear(); cur->id = id; cur->first_seq = seq; return true; };
        auto line = [&](const char *s, size_t len) {
            if (!len) return;                                                 // `if (read.empty()) continue;`
            if (!cur && !begin()) return;
            cur->add(s, len); ++seq;
            if (cur->n >= batch_reads) { work_q.push(cur); cur = nullptr; ++id; }
        };
        size_t got;
        while ((got = std::fread(buf.data(), 1, buf.size(), rf)) > 0 && !failed) {
            const char *p = buf.data(), *end = p + got;
            while (p < end) {
                const char *nl = (const char *) std::memchr(p, '\n', (size_t) (end - p));
                if (!nl) { carry.append(p, (size_t) (end - p)); break; }
                if (!carry.empty()) { carry.append(p, (size_t) (nl - p)); line(carry.data(), carry.size()); carry.clear(); }
                else line(p, (size_t) (nl - p));
                p = nl + 1;
            }
        }
        if (!carry.empty()) line(carry.data(), carry.size());
        if (cur && cur->n) { work_q.push(cur); ++id; }
        total_batches = id;
        work_q.close();
        out_cv.notify_all();
    });

    std::vector<std::thread> workers;
    for (int w = 0; w < n_workers; ++w) workers.emplace_back([&, w] {
        pgi_index *my = ix[w % gpus]; Batch *b;
        while (work_q.pop(b)) {
            pgi_result r; auto t0 = clk::now();
#ifdef PGI_QUERY_TAGS
            const int rc = pgi_query_tags(my, b->reads, b->off, b->n, nullptr, &r);
#else
            const int rc = pgi_find_mems(my, b->reads, b->off, b->n, mem_length, min_occ, nullptr, &r);
#endif
            if (rc != PGI_OK) {
                std::fprintf(stderr, "find_mems: batch %llu failed: %s\n", (unsigned long long) b->id, pgi_last_error());
                failed = true; out_cv.notify_all(); free_q.push(b); continue;
            }
            const double dt = std::chrono::duration<double>(clk::now() - t0).count();
            uint64_t len = 0;
#ifdef PGI_QUERY_TAGS
            char *txt = text ? pgi_format_query_tags_text(&r, b->first_seq - 1, &len) : nullptr;
#else
            char *txt = text ? pgi_format_text(&r, b->first_seq, 1, &len) : nullptr;
#endif
            { std::lock_guard<std::mutex> l(stat_mu); gpu_seconds += dt; total_reads += r.n_reads; total_mems += r.n_mems; total_tags += r.n_tags; }
            pgi_result_free(&r);
            const uint64_t id = b->id;
            free_q.push(b);
            { std::lock_guard<std::mutex> l(out_mu); done[id] = {txt, len}; }
            out_cv.notify_all();
        }
    });

    // ordered writer (this thread)
    auto t_run = clk::now();
    for (uint64_t next = 0;;) {
        std::pair<char *, uint64_t> item;
        {
            std::unique_lock<std::mutex> l(out_mu);
            out_cv.wait(l, [&] { return failed || done.count(next) || next >= total_batches.load(); });
            if (failed || !done.count(next)) break;
            item = done[next]; done.erase(next);
        }
        if (item.first) { std::fwrite(item.first, 1, item.second, stdout); pgi_free(item.first); }
        ++next;
    }
    reader.join(); free_q.close(); work_q.close();
    for (auto &t : workers) t.join();
    std::fclose(rf);
    const double wall = std::chrono::duration<double>(clk::now() - t_run).count();
    if (failed) return EXIT_FAILURE;
#ifndef PGI_QUERY_TAGS
    // src/find_mems.cpp:143-146 (values are not comparable: MEM finding and tag lookup are one fused GPU pass here)
    std::printf("\nTotal time for finding all MEMs: %g seconds\nTotal time for all tag queries: %g seconds\n", gpu_seconds, 0.0);
#ifdef PGI_CHUNKED      // src/find_mems_chunked.cpp:112-116
    std::printf("Chunked memory-mapped indices used: %g GB\nChunk size: %llu MB\n", image_gb, (unsigned long long) chunk_size_mb);
#elif defined(PGI_MMAP) // src/find_mems_memory_mapped.cpp:105-108
    std::printf("Memory-mapped indices used: %g GB\n", image_gb);
#endif
#else
    (void) mem_length; (void) min_occ; (void) gpu_seconds;
    if (total_reads == 0) { std::cerr << "No reads found in file." << std::endl; return 1; }     // src/query_tags.cpp:82-85
#endif
    std::fflush(stdout);
    std::cerr << "[find_mems] " << total_reads << " reads, " << total_mems << " MEMs, " << total_tags << " tags on " << gpus << " GPU(s) in " << wall
              << " s (" << (wall > 0 ? total_reads / wall : 0) << " reads/s including parsing and text output)" << std::endl;
    for (auto &b : pool) { pgi_host_free(b.reads); pgi_host_free(b.off); }
    for (auto *p : ix) pgi_index_free(p);
    return 0;
}
