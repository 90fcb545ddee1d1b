/* pgi.h — C ABI of the B200-native find_mems hot path (libpgi.so).
 *
 * This is the drop-in boundary. The reference (parsaeskandar/pangenome-index)
 * has no FFI layer of its own: its surface is the C++ classes panindexer::FastLocate / TagArray and
 * the header-only find_all_mems. Each entry point below names the reference interface it replaces
 * (file:line relative to the reference tree). Host C++ (the facade in pangenome-index_b200/host/)
 * and the Python test/bench harness call ONLY these functions; all CUDA lives behind them.
 *
 * Conventions: plain pointers and sizes, no C++ or torch types; every function returns 0 on success
 * or a negative pgi_status; the message of the last failure on the calling thread is available from
 * pgi_last_error(). No exception crosses this boundary. There is NO CPU fallback: every compute
 * entry point fails with PGI_ENODEVICE when no CUDA device is usable.
 */
#ifndef PGI_H
#define PGI_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    PGI_OK = 0,
    PGI_EINVAL = -1,     /* bad argument */
    PGI_EFORMAT = -2,    /* file is not a FastLocate .ri / tags file, or is corrupt / unsupported */
    PGI_ENODEVICE = -3,  /* no usable CUDA device */
    PGI_ECUDA = -4,      /* CUDA runtime error (message in pgi_last_error) */
    PGI_ENOMEM = -5
} pgi_status;

/* Tag file layouts. */
enum { PGI_TAGS_SDSL_COMPACT = 0,   /* TagArray::load_compressed_tags_compact, src/tag_arrays.cpp:766-776 */
       PGI_TAGS_BYTECODE_LEN = 1 }; /* TagArray::load_compressed_tags,         src/tag_arrays.cpp:739-763 */

/* Opaque device-resident index: the flattened r-index (+ optional tag arrays) on one GPU.
 * Immutable after creation; any number of host threads may query it concurrently. */
typedef struct pgi_index pgi_index;

/* Opaque reusable batch context: one CUDA stream + grow-only device / pinned-host work buffers. */
typedef struct pgi_batch pgi_batch;

/* struct MEM, include/pangenome_index/algorithm.hpp:644-651 (end is exclusive). */
typedef struct { uint64_t start, end, bwt_start; int64_t size; } pgi_mem;

/* FastLocate::bi_interval, include/pangenome_index/r-index.hpp:118-130. */
typedef struct { uint64_t forward, reverse; int64_t size; } pgi_bi_interval;

/* Result of one batch, in read order; MEMs of a read in discovery order.
 * Pointers are HOST pointers (pinned) for pgi_find_mems / pgi_batch_fetch, DEVICE pointers for
 * pgi_batch_device_result. They stay valid until the next run on the same batch or its destruction
 * (pgi_find_mems: until pgi_result_free). */
typedef struct {
    uint64_t n_reads, n_mems, n_tags;
    uint64_t *mem_off;   /* n_reads + 1 : MEMs of read i are mems[mem_off[i] .. mem_off[i+1])            */
    pgi_mem  *mems;      /* n_mems                                                                      */
    uint64_t *tag_off;   /* n_mems + 1  : tags of MEM m are tags[tag_off[m] .. tag_off[m+1])            */
    uint64_t *tags;      /* n_tags      : sorted unique encoded graph positions (node<<11|rev<<10|off)  */
    uint64_t *n_runs;    /* n_mems      : the reference's number_of_runs out-parameter                  */
    uint64_t extensions; /* FMD extensions of the reference algorithm (work counter; equals the oracle's E) */
    uint64_t seed_lookups; /* seed-table lookups; each stood for seed_k of those extensions (pgi_index_info_t) */
    uint64_t blocks_fetched; /* 64-byte run blocks the MEM kernel fetched from the index (work counter: the bytes it really moves) */
    void *owner;         /* internal */
} pgi_result;

/* Index facts (pgi_index_info). */
typedef struct {
    uint64_t n;              /* FastLocate::bwt_size(), r-index.hpp:337 */
    uint64_t sigma;          /* C.size() */
    uint64_t is_encoded;     /* file had ENCODED_BLOCKS (flag bit 0) */
    uint64_t n_ref_blocks;   /* ones in blocks_start_pos */
    uint64_t n_runs;         /* logical runs in the file */
    uint64_t n_dev_blocks;   /* 64-byte device run blocks (one per bucket of 2^bucket_shift BWT positions + chunk blocks) */
    uint64_t bucket_shift;   /* log2 of the positions covered by one bucket = one 64-byte block */
    uint64_t has_quirk_slot; /* legacy file with an absent nuc symbol */
    uint64_t has_tags, tag_layout, tag_runs, tag_items, tag_item_width, tag_universe;
    uint64_t device_bytes;   /* bytes of HBM held by the index */
    uint64_t n_crowded_buckets; /* buckets with more than 12 runs (directory block + chunk blocks: one more access) */
    uint64_t has_locate;     /* the file carries samples / last / last_to_run and pgi_locate_batch works on it */
    uint64_t max_length;     /* Header::max_length: SA value = seq_id * max_length + seq_offset (r-index.hpp:424-431) */
    uint64_t seed_k;         /* K of the seed table (4^K entries: the first K backward extensions from the full interval); 0 = none */
    uint64_t host_resident_bytes; /* out-of-core mode: bytes of the run-block array kept in pinned host memory and read over PCIe (0: all in HBM) */
    uint64_t C[8];
    uint8_t  sym_map[256];
} pgi_index_info_t;

const char *pgi_last_error(void);
int pgi_device_count(void);   /* number of CUDA devices (0 when none / no driver) */

/* Replaces FastLocate::load_encoded (src/r-index.cpp:406-459; falls back to the legacy layout of
 * FastLocate::load :378-404 when flag bit 0 is clear) + TagArray::load_compressed_tags_compact /
 * load_compressed_tags, and flattens both into the device layout on GPU `device`.
 * tag_bytes may be NULL (MEM finding only); ri_bytes may be NULL when tag_bytes is given (a tags-only index, the
 * TagArray facade: only pgi_tag_query_batch works on it). The input buffers are not retained. */
int pgi_index_create(const void *ri_bytes, size_t ri_len, const void *tag_bytes, size_t tag_len,
                     int tag_layout, int device, pgi_index **out);
/* GPU-side construction. Replaces, for an index that already exists as run-length arrays, the chain
 * FastLocate(rlbwt) constructor (block part, src/r-index.cpp:778-958) -> serialize_encoded (:297-376) -> load_encoded
 * (:406-459) and TagArray's compressed writers (src/tag_arrays.cpp:622-654, 940-974) -> load_compressed_tags_compact
 * (:766-776): the device index is built ON THE GPU from
 *   sym[i], len[i]          i < n_runs      the run-length BWT (symbol byte in {'\n','A','C','G','N','T'}, length > 0; what a
 *                                           .rl_bwt file holds, deps/grlBWT/include/bwt_io.h:40-58), in BWT order
 *   tag_val[j], tag_len[j]  j < n_tag_runs  the run-length tag array (encoded graph position node<<11|rev<<10|off, length > 0;
 *                                           runs longer than 511 are split as tag_arrays.cpp:942-957 does), or n_tag_runs = 0
 * device_pointers != 0: the four arrays are DEVICE pointers on GPU `device` (they are not modified and not retained).
 * The result equals pgi_index_create on the files the reference's writers emit for these runs (encoded .ri without
 * suffix-array samples -- so no pgi_locate_batch -- and sdsl-compact tags), answer for answer. No host copy of the index
 * is made. Limits of the device layout (both constructions): n < 2^40, fewer than 2^32 end markers and 'N's, fewer than
 * 2^32 - 1 run blocks of 2^bucket_shift BWT positions each (64 bytes per block: the index takes n / 2^bucket_shift * 64
 * bytes of HBM, bucket_shift <= 8, i.e. at least n / 4 bytes -- an index whose BWT is longer than about 4x the free HBM
 * does not fit one GPU; out-of-core operation is not built). */
int pgi_index_create_from_runs(const uint8_t *sym, const uint64_t *len, uint64_t n_runs,
                               const uint64_t *tag_val, const uint64_t *tag_len, uint64_t n_tag_runs,
                               int device_pointers, int device, pgi_index **out);
/* Replica of an index on another GPU (device-to-device copy over NVLink; no host round trip). */
int pgi_index_clone_to(const pgi_index *src, int device, pgi_index **out);
void pgi_index_free(pgi_index *);
int pgi_index_info(const pgi_index *, pgi_index_info_t *out);

/* ---- Cached flattened index image.
 * Replaces MemoryMappedIndexManager / ChunkedMemoryMappedIndexManager::open (src/find_mems_memory_mapped.cpp:36-45,
 * src/find_mems_chunked.cpp:29-50: the .ri / .tags files are mmap'ed, whole or in chunks of chunk_size_mb, instead of
 * being deserialised). Here the SDSL files are flattened once into the device layout and written as ONE image file;
 * opening it maps the file read-only and streams it to HBM through two pinned staging buffers of chunk_bytes: no
 * parsing, no host copy of the index, host memory bounded by 2 x chunk_bytes however large the index. */
typedef struct {
    uint64_t file_bytes, payload_bytes;
    uint64_t n, sigma, n_dev_blocks, bucket_shift, is_encoded;
    uint64_t has_ri, has_tags, tag_layout, has_locate;
    uint64_t src_ri_bytes, src_tag_bytes;   /* sizes of the files the image was made from                              */
    uint64_t src_ri_sum, src_tag_sum;       /* pgi_image_source_sum of those files (staleness check: whole-file checksum) */
    uint64_t chunk_bytes, chunks_copied;    /* filled by pgi_index_open_image only                                      */
} pgi_image_info_t;
/* Host-only: flatten the files (same arguments as pgi_index_create) and write the image to `path` (atomically:
 * path.tmp + rename). Needs no CUDA device. */
int pgi_image_write(const void *ri_bytes, size_t ri_len, const void *tag_bytes, size_t tag_len, int tag_layout, const char *path);
/* Host-only: validate the header of an image (verify != 0: also the checksum of every section) and describe it. */
int pgi_image_info(const char *path, int verify, pgi_image_info_t *out);
/* Content stamp of a source file (FNV-1a over the whole file, near memory speed); compared with src_*_sum. */
uint64_t pgi_image_source_sum(const void *bytes, size_t len);
/* Map an image and upload it to GPU `device`. chunk_bytes = 0: 64 MiB (the reference's chunk_size_mb argument, in
 * bytes). verify != 0 (or PGI_IMAGE_VERIFY=1): check every section checksum first. The resulting index is
 * indistinguishable from one made by pgi_index_create from the same files. `info` may be NULL. */
int pgi_index_open_image(const char *path, int device, uint64_t chunk_bytes, int verify, pgi_index **out, pgi_image_info_t *info);

/* Batched FastLocate::backward_extend_encoded (src/r-index.cpp:713-756) / forward_extend_encoded
 * (:758-764): out[i] = extend(in[i], sym[i]); dir[i] 0 = backward, 1 = forward. Host buffers. */
int pgi_extend_batch(pgi_index *, uint64_t n, const pgi_bi_interval *in, const uint8_t *sym,
                     const uint8_t *dir, pgi_bi_interval *out);
/* Batched FastLocate::rank_at_cached_encoded (src/r-index.cpp:619-641): out8[8*i + j] = R(pos[i])[j], j < sigma. */
int pgi_rank_batch(pgi_index *, uint64_t n, const uint64_t *pos, uint64_t *out8);
/* pgi_rank_batch without any device allocation. No counterpart in the reference: added
 * for callers that may not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it). Same arguments and results as pgi_rank_batch (host buffers), but the device staging of positions and
 * rank vectors is a static __device__ array of 72 * PGI_RANK_NOALLOC_MAX bytes (4.5 MiB) that every GPU loading libpgi.so
 * holds until the process exits; the call issues no cudaMalloc / cudaFree / allocating library routine.
 * n > PGI_RANK_NOALLOC_MAX fails with PGI_EINVAL (split the batch, or call pgi_rank_batch). All calls share the one
 * array: calls are serialised inside a process by a mutex and complete before returning; work of other origin that
 * would touch the array while a call is on the GPU races with it. */
#define PGI_RANK_NOALLOC_MAX 65536
int pgi_rank_batch_noalloc(pgi_index *, uint64_t n, const uint64_t *pos, uint64_t *out8);
/* Batched TagArray::query_compressed_compact (src/tag_arrays.cpp:856-890) / query_compressed (:780-854),
 * chosen by the layout the index was created with. Intervals are [start[i], end[i]] inclusive.
 * Result: tag_off / tags / n_runs of `out` (n_mems = n; mems, mem_off are NULL). Free with pgi_result_free. */
int pgi_tag_query_batch(pgi_index *, uint64_t n, const uint64_t *start, const uint64_t *end, pgi_result *out);

/* Batched FastLocate::locate_encoded (src/r-index.cpp:1300-1343): for every BWT range [first[i],
 * second[i]] (inclusive; second < first = empty) the SA value of its first row from the sample of that row's run
 * (getSample, include/pangenome_index/r-index.hpp:616) plus one phi step (locateNext, src/r-index.cpp:1369-1372) per row.
 * Result (malloc'd host arrays, free with pgi_locate_result_free):
 *   ids[id_off[i] .. id_off[i+1])  the reference's return value: sorted unique SEQUENCE IDS (seqId, r-index.hpp:429)
 *   sa[row_off[i] + t]             the SA value (seq_id * max_length + seq_offset) of row first[i] + t, i.e. the values
 *                                  the reference holds in `first` while it walks the range (decompressSA, :1345-1356)
 * The reference's legacy twin FastLocate::locate has an empty body (:1281-1296); on a legacy file the same algorithm runs
 * over the file's own logical runs. PGI_EFORMAT when the file has no samples or is an encoded file without N (the
 * reference's run lookup skips six cumulatives where five exist); PGI_EINVAL for a range beyond the BWT. */
typedef struct {
    uint64_t n, n_ids, n_rows;
    uint64_t *id_off, *ids;     /* n + 1, n_ids  */
    uint64_t *row_off, *sa;     /* n + 1, n_rows */
} pgi_locate_result;
int pgi_locate_batch(pgi_index *, uint64_t n, const uint64_t *first, const uint64_t *second, pgi_locate_result *out);
void pgi_locate_result_free(pgi_locate_result *);

/* The hot path: for every read, find_all_mems (include/pangenome_index/algorithm.hpp:739-757) and,
 * when the index has tags, the tag query of src/find_mems.cpp:128 for every MEM.
 * reads = concatenated read bytes, read i = reads[read_off[i] .. read_off[i+1]). HOST buffers (pinned
 * recommended); host->device and device->host copies happen inside. cuda_stream: a cudaStream_t or NULL
 * (library-owned stream). Blocks until the result is complete in host memory. Free with pgi_result_free. */
int pgi_find_mems(pgi_index *, const char *reads, const uint64_t *read_off, uint64_t n_reads,
                  uint64_t min_len, uint64_t min_occ, void *cuda_stream, pgi_result *out);
void pgi_result_free(pgi_result *);

/* ---- Compact wire form of a find_mems result: the same information as pgi_result in fewer bytes across PCIe (a batch of
 * 2 M reads of 150 bp: 117 MB instead of 227 MB), packed on the GPU and expanded on the host only when asked to.
 *   mem_cnt[i]            MEMs of read i (u32)                       -> pgi_result.mem_off by prefix sum
 *   rec[2 m], rec[2 m+1]  bwt_start | start << 40,  size | (end - start) << 40    (positions < 2^40, reads < 2^24 bases)
 *   runs_tags[2 m], [+1]  number_of_runs, number of unique tags of MEM m (u32 each)  -> n_runs, tag_off
 *   tags                  tag_bytes (5 or 8) little-endian bytes per tag, all MEMs back to back
 * pgi_find_mems_compact fails with PGI_EINVAL when a value does not fit (a read of 2^24 bases or more, a count of 2^32 or
 * more); use pgi_find_mems then. Pointers are pinned host memory owned by the library until pgi_result_compact_free. */
typedef struct {
    uint64_t n_reads, n_mems, n_tags, tag_bytes;
    uint32_t *mem_cnt; uint64_t *rec; uint32_t *runs_tags; uint8_t *tags;
    uint64_t extensions, seed_lookups, blocks_fetched;
    void *owner;
} pgi_result_compact;
/* pgi_find_mems with the compact result. read_off == NULL: every read has read_len bytes (no offset array crosses PCIe). */
int pgi_find_mems_compact(pgi_index *, const char *reads, const uint64_t *read_off, uint64_t read_len, uint64_t n_reads,
                          uint64_t min_len, uint64_t min_occ, void *cuda_stream, pgi_result_compact *out);
void pgi_result_compact_free(pgi_result_compact *);
/* Host-side expansion into the reference-layout arrays (malloc'd; free with pgi_result_expanded_free). */
int pgi_result_expand(const pgi_result_compact *, pgi_result *out);
void pgi_result_expanded_free(pgi_result *);
/* Batch-context forms: stage fixed-length reads without an offset array; pack + copy the last run's result. */
int pgi_batch_upload_fixed(pgi_batch *, const char *reads, uint64_t n_reads, uint64_t read_len);
int pgi_batch_fetch_compact(pgi_batch *, pgi_result_compact *out);

/* The query_tags path: for every read the whole-read backward search FastLocate::count_encoded
 * (include/pangenome_index/r-index.hpp:550-556; the legacy FastLocate::count :541-547 when the file has no encoded
 * blocks, as src/query_tags.cpp:90-96 chooses) and, for reads with a match and an index with tags, the tag query of
 * the final range (src/query_tags.cpp:97-102). Same buffer conventions as pgi_find_mems. Result: ONE record per read,
 *   mems[i] = { start 0, end = read length, bwt_start = range.first, size = range.second - range.first + 1 },
 * an empty range -- the reference's {1,0} -- being bwt_start 1, size 0 (no tags, n_runs 0); mem_off[i] = i;
 * n_runs[i] = the reference's number_of_runs; extensions = LF steps executed. An empty read keeps the full range.
 * PGI_EFORMAT on an encoded index without N (sigma 5): the reference reads a run header as a sixth cumulative there
 * (src/r-index.cpp:577) and its results are garbage that this library does not reproduce. */
int pgi_query_tags(pgi_index *, const char *reads, const uint64_t *read_off, uint64_t n_reads,
                   void *cuda_stream, pgi_result *out);
/* The same on a batch context (after pgi_batch_upload / pgi_batch_set_device_reads); results via
 * pgi_batch_device_result / pgi_batch_fetch like pgi_batch_run. */
int pgi_batch_run_query_tags(pgi_batch *);
/* stdout of src/query_tags.cpp for such a result (the two lines TagArray::query_compressed_compact prints itself,
 * src/tag_arrays.cpp:885-889, then the read_index summary line :103-108); read_index counts from first_index.
 * malloc'd, free with pgi_free. */
char *pgi_format_query_tags_text(const pgi_result *, uint64_t first_index, uint64_t *len);

/* Explicit batch contexts, for callers that pipeline several batches per GPU (the find_mems driver)
 * or keep inputs / results resident in HBM (bench.py's device-resident figure). */
int pgi_batch_create(pgi_index *, void *cuda_stream /* or NULL */, pgi_batch **out);
void pgi_batch_free(pgi_batch *);
/* Asynchronously stage host reads into the batch's device buffers (pinned source recommended). */
int pgi_batch_upload(pgi_batch *, const char *reads, const uint64_t *read_off, uint64_t n_reads);
/* Use reads already resident in device memory (d_reads: 8-byte aligned and readable up to the next multiple of 8 bytes;
 * total_bytes = d_read_off[n_reads], the byte length of the reads -- the seed pre-pass packs exactly that many bytes). */
int pgi_batch_set_device_reads(pgi_batch *, const char *d_reads, const uint64_t *d_read_off,
                               uint64_t n_reads, uint64_t total_bytes);
/* Run MEM finding + tag lookup on the staged reads (the body of the read loop, src/find_mems.cpp:94-139, for the whole
 * batch); results stay in HBM. Returns after the last kernel has completed. */
int pgi_batch_run(pgi_batch *, uint64_t min_len, uint64_t min_occ);
/* The same step in two halves, for callers that keep several batches in flight on one GPU from one host thread:
 * pgi_batch_run_async queues every kernel of the step on the batch's stream and returns at once (the step is one
 * stream-ordered sequence; nothing on the host is waited for); pgi_batch_wait blocks until the step has finished, fetches
 * its counters and -- rarely -- runs a step again whose buffers did not fit. pgi_batch_run = the two back to back. Steps
 * of different batches (different streams) overlap on the GPU: the next step's persistent MEM kernel fills the SMs that the
 * previous one's last reads leave idle. The reads of a batch must stay untouched between the two calls. */
int pgi_batch_run_async(pgi_batch *, uint64_t min_len, uint64_t min_occ);
int pgi_batch_wait(pgi_batch *);
/* Device pointers of the last run's result. */
int pgi_batch_device_result(pgi_batch *, pgi_result *out);
/* Copy the last run's result into the batch's pinned host buffers and wait for it. */
int pgi_batch_fetch(pgi_batch *, pgi_result *out);
/* Device time of the stages of the last run, from CUDA events on the batch's stream, in milliseconds:
 * [0] MEM kernel (fmd_mem_kernel), [1] MEM ordering (scan + scatter), [2] tag sizing, [3] tag decode + sort,
 * [4] tag unique + compaction, [5] whole run, [6] number of this library's own kernel launches in the run (CUB scans / sorts excluded), [7] MEM kernel re-runs. */
int pgi_batch_stage_ms(pgi_batch *, double ms[8]);

/* Text of src/find_mems.cpp:115-139 (+ tag lines of src/tag_arrays.cpp:885-889) for a host result;
 * sequence numbers start at first_seq. Host-side formatting (no GPU work). Returns a malloc'd buffer
 * (free with pgi_free), length in *len. with_tags = 0 omits the two tag lines per MEM. */
/* Pinned (page-locked) host memory for read / offset staging buffers: copies from it run at full PCIe rate and
 * asynchronously. Returns NULL on failure. Free with pgi_host_free. */
void *pgi_host_alloc(size_t bytes);
void pgi_host_free(void *);

char *pgi_format_text(const pgi_result *, uint64_t first_seq, int with_tags, uint64_t *len);
void pgi_free(void *);

/* M1 microbenchmark: independent random `bytes_per_access`-byte (32, 64 or 128) gathers
 * over a buffer of buf_bytes on `device`; returns GB/s of useful bytes in *gbs. The random-access roofline. */
int pgi_random_access_gbs(int device, uint64_t buf_bytes, uint32_t bytes_per_access, uint64_t n_access,
                          int chain /* >= 2: dependent chain of that length per thread; 1: four independent loads in flight per
                                       thread; 0: the CEILING -- teams of bytes_per_access / 16 lanes per slot, eight independent
                                       slots in flight per lane, full occupancy (the MEM kernel's fetch pattern, no arithmetic) */,
                          double *gbs);

/* cudaLimitMaxL2FetchGranularity of `device` (bytes fetched from DRAM around a missing 32-byte sector: 32, 64 or 128).
 * set_bytes != 0 sets it first; *current receives the value in effect. A device-wide setting of the calling process. */
int pgi_l2_fetch_granularity(int device, int set_bytes, int *current);

#ifdef __cplusplus
}
#endif
#endif
