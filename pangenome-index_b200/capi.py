"""ctypes binding of include/pgi.h (libpgi.so). Plumbing only: every call lands in the CUDA library.
If the library is missing this module raises — there is no Python / CPU fallback for the query path."""
import ctypes, os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PGI_LIB") or os.path.join(_HERE, "libpgi.so")      # PGI_LIB: the bounds-checking build (csrc/Makefile)
_lib = None
RANK_NOALLOC_MAX = 65536      # PGI_RANK_NOALLOC_MAX of include/pgi.h

SYMBOLS = ["pgi_last_error", "pgi_device_count", "pgi_index_create", "pgi_index_clone_to", "pgi_index_free", "pgi_index_info",
           "pgi_extend_batch", "pgi_rank_batch", "pgi_tag_query_batch", "pgi_find_mems", "pgi_result_free", "pgi_batch_create",
           "pgi_batch_free", "pgi_batch_upload", "pgi_batch_set_device_reads", "pgi_batch_run", "pgi_batch_device_result",
           "pgi_batch_fetch", "pgi_batch_stage_ms", "pgi_format_text", "pgi_free", "pgi_query_tags", "pgi_batch_run_query_tags", "pgi_format_query_tags_text", "pgi_locate_batch", "pgi_locate_result_free", "pgi_random_access_gbs", "pgi_host_alloc", "pgi_host_free",
           "pgi_image_write", "pgi_image_info", "pgi_image_source_sum", "pgi_index_open_image", "pgi_index_create_from_runs", "pgi_l2_fetch_granularity", "pgi_find_mems_compact", "pgi_result_compact_free",
           "pgi_result_expand", "pgi_result_expanded_free", "pgi_batch_upload_fixed", "pgi_batch_fetch_compact", "pgi_rank_batch_noalloc", "pgi_batch_run_async", "pgi_batch_wait"]

class Result(ctypes.Structure):
    _fields_ = [("n_reads", ctypes.c_uint64), ("n_mems", ctypes.c_uint64), ("n_tags", ctypes.c_uint64),
                ("mem_off", ctypes.c_void_p), ("mems", ctypes.c_void_p), ("tag_off", ctypes.c_void_p),
                ("tags", ctypes.c_void_p), ("n_runs", ctypes.c_void_p), ("extensions", ctypes.c_uint64), ("seed_lookups", ctypes.c_uint64), ("blocks_fetched", ctypes.c_uint64), ("owner", ctypes.c_void_p)]

class ResultCompact(ctypes.Structure):
    _fields_ = [("n_reads", ctypes.c_uint64), ("n_mems", ctypes.c_uint64), ("n_tags", ctypes.c_uint64), ("tag_bytes", ctypes.c_uint64),
                ("mem_cnt", ctypes.c_void_p), ("rec", ctypes.c_void_p), ("runs_tags", ctypes.c_void_p), ("tags", ctypes.c_void_p),
                ("extensions", ctypes.c_uint64), ("seed_lookups", ctypes.c_uint64), ("blocks_fetched", ctypes.c_uint64), ("owner", ctypes.c_void_p)]

class IndexInfo(ctypes.Structure):
    _fields_ = [(k, ctypes.c_uint64) for k in ["n", "sigma", "is_encoded", "n_ref_blocks", "n_runs", "n_dev_blocks", "bucket_shift",
                                               "has_quirk_slot", "has_tags", "tag_layout", "tag_runs", "tag_items", "tag_item_width",
                                               "tag_universe", "device_bytes", "n_crowded_buckets", "has_locate", "max_length", "seed_k", "host_resident_bytes"]] + [("C", ctypes.c_uint64 * 8), ("sym_map", ctypes.c_uint8 * 256)]

class LocateResult(ctypes.Structure):
    _fields_ = [("n", ctypes.c_uint64), ("n_ids", ctypes.c_uint64), ("n_rows", ctypes.c_uint64), ("id_off", ctypes.c_void_p),
                ("ids", ctypes.c_void_p), ("row_off", ctypes.c_void_p), ("sa", ctypes.c_void_p)]

class ImageInfo(ctypes.Structure):
    _fields_ = [(k, ctypes.c_uint64) for k in ["file_bytes", "payload_bytes", "n", "sigma", "n_dev_blocks", "bucket_shift", "is_encoded", "has_ri",
                                               "has_tags", "tag_layout", "has_locate", "src_ri_bytes", "src_tag_bytes", "src_ri_sum", "src_tag_sum",
                                               "chunk_bytes", "chunks_copied"]]

class PgiError(RuntimeError):
    pass

def build():
    import subprocess
    subprocess.check_call(["make", "-s", "-C", os.path.join(_HERE, "csrc")])

def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PgiError("libpgi.so is not built (make -C pangenome-index_b200/csrc); there is no CPU fallback")
        L = ctypes.CDLL(LIB_PATH)
        L.pgi_last_error.restype = ctypes.c_char_p
        L.pgi_format_text.restype = ctypes.c_void_p
        vp, u64, i32 = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int
        L.pgi_index_create.argtypes = [vp, ctypes.c_size_t, vp, ctypes.c_size_t, i32, i32, ctypes.POINTER(vp)]
        L.pgi_index_clone_to.argtypes = [vp, i32, ctypes.POINTER(vp)]
        L.pgi_find_mems_compact.argtypes = [vp, vp, vp, u64, u64, u64, u64, vp, ctypes.POINTER(ResultCompact)]
        L.pgi_result_compact_free.argtypes = [ctypes.POINTER(ResultCompact)]
        L.pgi_result_expand.argtypes = [ctypes.POINTER(ResultCompact), ctypes.POINTER(Result)]
        L.pgi_result_expanded_free.argtypes = [ctypes.POINTER(Result)]
        L.pgi_batch_upload_fixed.argtypes = [vp, vp, u64, u64]
        L.pgi_batch_fetch_compact.argtypes = [vp, ctypes.POINTER(ResultCompact)]
        L.pgi_index_create_from_runs.argtypes = [vp, vp, u64, vp, vp, u64, i32, i32, ctypes.POINTER(vp)]
        L.pgi_index_free.argtypes = [vp]
        L.pgi_index_info.argtypes = [vp, ctypes.POINTER(IndexInfo)]
        L.pgi_extend_batch.argtypes = [vp, u64, vp, vp, vp, vp]
        L.pgi_rank_batch.argtypes = [vp, u64, vp, vp]
        L.pgi_rank_batch_noalloc.argtypes = [vp, u64, vp, vp]
        L.pgi_tag_query_batch.argtypes = [vp, u64, vp, vp, ctypes.POINTER(Result)]
        L.pgi_find_mems.argtypes = [vp, vp, vp, u64, u64, u64, vp, ctypes.POINTER(Result)]
        L.pgi_result_free.argtypes = [ctypes.POINTER(Result)]
        L.pgi_batch_create.argtypes = [vp, vp, ctypes.POINTER(vp)]
        L.pgi_batch_free.argtypes = [vp]
        L.pgi_batch_upload.argtypes = [vp, vp, vp, u64]
        L.pgi_batch_set_device_reads.argtypes = [vp, vp, vp, u64, u64]
        L.pgi_batch_run.argtypes = [vp, u64, u64]
        L.pgi_batch_run_async.argtypes = [vp, u64, u64]
        L.pgi_batch_wait.argtypes = [vp]
        L.pgi_batch_device_result.argtypes = [vp, ctypes.POINTER(Result)]
        L.pgi_batch_fetch.argtypes = [vp, ctypes.POINTER(Result)]
        L.pgi_batch_stage_ms.argtypes = [vp, ctypes.POINTER(ctypes.c_double * 8)]
        L.pgi_format_text.argtypes = [ctypes.POINTER(Result), u64, i32, ctypes.POINTER(u64)]
        L.pgi_free.argtypes = [vp]
        L.pgi_query_tags.argtypes = [vp, vp, vp, u64, vp, ctypes.POINTER(Result)]
        L.pgi_batch_run_query_tags.argtypes = [vp]
        L.pgi_locate_batch.argtypes = [vp, u64, vp, vp, ctypes.POINTER(LocateResult)]
        L.pgi_locate_result_free.argtypes = [ctypes.POINTER(LocateResult)]
        L.pgi_format_query_tags_text.restype = ctypes.c_void_p
        L.pgi_format_query_tags_text.argtypes = [ctypes.POINTER(Result), u64, ctypes.POINTER(u64)]
        L.pgi_host_alloc.restype = vp; L.pgi_host_alloc.argtypes = [ctypes.c_size_t]; L.pgi_host_free.argtypes = [vp]
        L.pgi_random_access_gbs.argtypes = [i32, u64, ctypes.c_uint32, u64, i32, ctypes.POINTER(ctypes.c_double)]
        L.pgi_image_write.argtypes = [vp, ctypes.c_size_t, vp, ctypes.c_size_t, i32, ctypes.c_char_p]
        L.pgi_image_info.argtypes = [ctypes.c_char_p, i32, ctypes.POINTER(ImageInfo)]
        L.pgi_image_source_sum.restype = u64; L.pgi_image_source_sum.argtypes = [vp, ctypes.c_size_t]
        L.pgi_index_open_image.argtypes = [ctypes.c_char_p, i32, u64, i32, ctypes.POINTER(vp), ctypes.POINTER(ImageInfo)]
        _lib = L
    return _lib

def _ck(rc):
    if rc != 0:
        raise PgiError("pgi error %d: %s" % (rc, lib().pgi_last_error().decode()))

def _arr(ptr, n, dtype=np.uint64):
    if not ptr or n == 0:
        return np.zeros(0, dtype=dtype)
    buf = (ctypes.c_uint8 * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).copy()

def _result_to_dict(res, L, with_tags, text=True):
    out = dict(n_reads=res.n_reads, n_mems=res.n_mems, n_tags=res.n_tags, extensions=res.extensions, seed_lookups=res.seed_lookups, blocks_fetched=res.blocks_fetched,
               mem_off=_arr(res.mem_off, res.n_reads + 1 if res.mem_off else 0),
               mems=_arr(res.mems, 4 * res.n_mems if res.mems else 0).reshape(-1, 4),
               tag_off=_arr(res.tag_off, res.n_mems + 1), tags=_arr(res.tags, res.n_tags), n_runs=_arr(res.n_runs, res.n_mems))
    if text and res.mem_off:
        ln = ctypes.c_uint64()
        p = L.pgi_format_text(ctypes.byref(res), 1, int(with_tags), ctypes.byref(ln))
        out["text"] = ctypes.string_at(p, ln.value); L.pgi_free(p)
    return out

def image_write(ri: bytes, tags: bytes, tag_layout, path):
    """Flatten the SDSL files on the host and write the cached device image (no GPU needed)."""
    _ck(lib().pgi_image_write(ri, len(ri) if ri else 0, tags, len(tags) if tags else 0, tag_layout, os.fsencode(path)))

def image_info(path, verify=False):
    f = ImageInfo(); _ck(lib().pgi_image_info(os.fsencode(path), int(verify), ctypes.byref(f)))
    return {k: getattr(f, k) for k, _ in ImageInfo._fields_}

def image_source_sum(data: bytes):
    return lib().pgi_image_source_sum(data, len(data))

class Index:
    """Device-resident index (FastLocate + optional TagArray) on one GPU."""
    @classmethod
    def open_image(cls, path, device=0, chunk_bytes=0, verify=False):
        """mmap a cached image and stream it to HBM in chunks (the memory-mapped / chunked index managers of the reference)."""
        h = ctypes.c_void_p(); f = ImageInfo()
        _ck(lib().pgi_index_open_image(os.fsencode(path), device, chunk_bytes, int(verify), ctypes.byref(h), ctypes.byref(f)))
        ix = cls(None, device=device, _handle=h); ix.image = {k: getattr(f, k) for k, _ in ImageInfo._fields_}
        return ix
    @classmethod
    def from_runs(cls, sym, length, tag_val=None, tag_len=None, device=0):
        """GPU-side construction from the run-length BWT (+ run-length tags). numpy arrays (host) or objects with
        .data_ptr() (torch CUDA tensors: uint8 / int64 / int64 / int64 on `device`)."""
        h = ctypes.c_void_p(); dev_ptrs = hasattr(sym, "data_ptr")
        if dev_ptrs:
            import torch
            assert sym.dtype == torch.uint8 and length.dtype == torch.int64 and sym.is_contiguous() and length.is_contiguous()
            a, b, n = sym.data_ptr(), length.data_ptr(), sym.numel()
            c = tag_val.data_ptr() if tag_val is not None else None; d = tag_len.data_ptr() if tag_val is not None else None
            nt = tag_val.numel() if tag_val is not None else 0
        else:
            sym = np.ascontiguousarray(sym, dtype=np.uint8); length = np.ascontiguousarray(length, dtype=np.uint64)
            a, b, n = sym.ctypes.data, length.ctypes.data, len(sym); c = d = None; nt = 0
            if tag_val is not None:
                tag_val = np.ascontiguousarray(tag_val, dtype=np.uint64); tag_len = np.ascontiguousarray(tag_len, dtype=np.uint64)
                c, d, nt = tag_val.ctypes.data, tag_len.ctypes.data, len(tag_val)
        _ck(lib().pgi_index_create_from_runs(a, b, n, c, d, nt, int(dev_ptrs), device, ctypes.byref(h)))
        return cls(None, device=device, _handle=h)
    def __init__(self, ri: bytes, tags: bytes = None, tag_layout=0, device=0, _handle=None):
        L = lib(); self.device = device
        if _handle is not None:
            self.h = _handle
        else:
            h = ctypes.c_void_p()
            _ck(L.pgi_index_create(ri, len(ri) if ri else 0, tags, len(tags) if tags else 0, tag_layout, device, ctypes.byref(h)))
            self.h = h
        info = IndexInfo(); _ck(L.pgi_index_info(self.h, ctypes.byref(info)))
        self.info = {k: getattr(info, k) for k, _ in IndexInfo._fields_[:20]}
        self.n = info.n; self.sigma = info.sigma; self.has_tags = bool(info.has_tags)
        self.C = list(info.C)[: info.sigma]; self.sym_map = bytes(info.sym_map)
    def clone_to(self, device):
        h = ctypes.c_void_p(); _ck(lib().pgi_index_clone_to(self.h, device, ctypes.byref(h)))
        return Index(None, device=device, _handle=h)
    def close(self):
        if getattr(self, "h", None):
            lib().pgi_index_free(self.h); self.h = None
    def __del__(self):
        try: self.close()
        except Exception: pass
    def extend_batch(self, iv3, sym, direction):
        iv3 = np.ascontiguousarray(iv3, dtype=np.uint64); sym = np.ascontiguousarray(sym, dtype=np.uint8)
        direction = np.ascontiguousarray(direction, dtype=np.uint8); out = np.zeros_like(iv3)
        _ck(lib().pgi_extend_batch(self.h, len(sym), iv3.ctypes.data, sym.ctypes.data, direction.ctypes.data, out.ctypes.data))
        return out
    def rank_batch(self, pos):
        pos = np.ascontiguousarray(pos, dtype=np.uint64); out = np.zeros((len(pos), 8), dtype=np.uint64)
        _ck(lib().pgi_rank_batch(self.h, len(pos), pos.ctypes.data, out.ctypes.data))
        return out
    def rank_batch_noalloc(self, pos):
        """pgi_rank_batch_noalloc: static device scratch, at most RANK_NOALLOC_MAX positions per call."""
        pos = np.ascontiguousarray(pos, dtype=np.uint64); out = np.zeros((len(pos), 8), dtype=np.uint64)
        _ck(lib().pgi_rank_batch_noalloc(self.h, len(pos), pos.ctypes.data, out.ctypes.data))
        return out
    def tag_query_batch(self, start, end):
        start = np.ascontiguousarray(start, dtype=np.uint64); end = np.ascontiguousarray(end, dtype=np.uint64)
        res = Result(); L = lib()
        _ck(L.pgi_tag_query_batch(self.h, len(start), start.ctypes.data, end.ctypes.data, ctypes.byref(res)))
        out = _result_to_dict(res, L, True, text=False); L.pgi_result_free(ctypes.byref(res))
        return out
    def find_mems(self, reads, read_off, min_len, min_occ, text=True):
        """reads: uint8 concatenation (host), read_off: uint64 [n+1]. Host buffers in, host arrays out (the e2e path)."""
        reads = np.ascontiguousarray(reads, dtype=np.uint8); read_off = np.ascontiguousarray(read_off, dtype=np.uint64)
        res = Result(); L = lib()
        _ck(L.pgi_find_mems(self.h, reads.ctypes.data, read_off.ctypes.data, len(read_off) - 1, min_len, min_occ, None, ctypes.byref(res)))
        out = _result_to_dict(res, L, self.has_tags, text=text); L.pgi_result_free(ctypes.byref(res))
        return out

    def query_tags(self, reads, read_off, text=True):
        """The query_tags path (src/query_tags.cpp:87-109): whole-read count + tag query; one record per read."""
        reads = np.ascontiguousarray(reads, dtype=np.uint8); read_off = np.ascontiguousarray(read_off, dtype=np.uint64)
        res = Result(); L = lib()
        _ck(L.pgi_query_tags(self.h, reads.ctypes.data, read_off.ctypes.data, len(read_off) - 1, None, ctypes.byref(res)))
        out = _result_to_dict(res, L, self.has_tags, text=False)
        if text:
            ln = ctypes.c_uint64()
            p = L.pgi_format_query_tags_text(ctypes.byref(res), 0, ctypes.byref(ln))
            out["text"] = ctypes.string_at(p, ln.value); L.pgi_free(p)
        L.pgi_result_free(ctypes.byref(res))
        return out

    def locate_batch(self, first, second):
        """FastLocate::locate_encoded per BWT range -> (id_off [n+1], sorted unique sequence ids, row_off [n+1], SA values row by row)."""
        first = np.ascontiguousarray(first, dtype=np.uint64); second = np.ascontiguousarray(second, dtype=np.uint64)
        res = LocateResult(); L = lib()
        _ck(L.pgi_locate_batch(self.h, len(first), first.ctypes.data, second.ctypes.data, ctypes.byref(res)))
        out = (_arr(res.id_off, res.n + 1), _arr(res.ids, res.n_ids), _arr(res.row_off, res.n + 1), _arr(res.sa, res.n_rows))
        L.pgi_locate_result_free(ctypes.byref(res))
        return out

    def find_mems_raw(self, reads_ptr, off_ptr, n_reads, min_len, min_occ):
        """pgi_find_mems on raw host pointers (pinned recommended); returns the Result struct whose arrays live in
        library-owned pinned memory until result_free(res). No numpy copies: this is what bench.py times end to end."""
        res = Result()
        _ck(lib().pgi_find_mems(self.h, reads_ptr, off_ptr, n_reads, min_len, min_occ, None, ctypes.byref(res)))
        return res
    @staticmethod
    def result_free(res):
        lib().pgi_result_free(ctypes.byref(res))
    def find_mems_compact_raw(self, reads_ptr, off_ptr, read_len, n_reads, min_len, min_occ):
        """pgi_find_mems_compact on raw host pointers (off_ptr None: fixed-length reads); free with result_compact_free."""
        res = ResultCompact()
        _ck(lib().pgi_find_mems_compact(self.h, reads_ptr, off_ptr, read_len, n_reads, min_len, min_occ, None, ctypes.byref(res)))
        return res
    @staticmethod
    def result_compact_free(res):
        lib().pgi_result_compact_free(ctypes.byref(res))
    def find_mems_compact(self, reads, read_len, min_len, min_occ, read_off=None):
        """Compact call + host expansion -> the same dict as find_mems (text=False)."""
        reads = np.ascontiguousarray(reads, dtype=np.uint8); L = lib()
        if read_off is not None: read_off = np.ascontiguousarray(read_off, dtype=np.uint64)
        n = (len(read_off) - 1) if read_off is not None else (len(reads) // read_len if read_len else 0)
        c = self.find_mems_compact_raw(reads.ctypes.data, read_off.ctypes.data if read_off is not None else None, read_len, n, min_len, min_occ)
        wire = c.n_reads * 4 + c.n_mems * 24 + c.n_tags * c.tag_bytes
        r = Result(); _ck(L.pgi_result_expand(ctypes.byref(c), ctypes.byref(r)))
        out = _result_to_dict(r, L, self.has_tags, text=False); out["wire_bytes"] = wire
        L.pgi_result_expanded_free(ctypes.byref(r)); self.result_compact_free(c)
        return out

class Batch:
    """Reusable batch context (stream + work buffers) for pipelined or device-resident runs."""
    def __init__(self, index: Index, stream=None):
        self.index = index; h = ctypes.c_void_p()
        _ck(lib().pgi_batch_create(index.h, stream, ctypes.byref(h))); self.h = h
    def close(self):
        if getattr(self, "h", None):
            lib().pgi_batch_free(self.h); self.h = None
    def __del__(self):
        try: self.close()
        except Exception: pass
    def upload(self, reads_ptr, off_ptr, n_reads):
        _ck(lib().pgi_batch_upload(self.h, reads_ptr, off_ptr, n_reads))
    def set_device_reads(self, d_reads_ptr, d_off_ptr, n_reads, total_bytes):
        _ck(lib().pgi_batch_set_device_reads(self.h, d_reads_ptr, d_off_ptr, n_reads, total_bytes))
    def run(self, min_len, min_occ):
        _ck(lib().pgi_batch_run(self.h, min_len, min_occ))
    def run_async(self, min_len, min_occ):
        _ck(lib().pgi_batch_run_async(self.h, min_len, min_occ))
    def wait(self):
        _ck(lib().pgi_batch_wait(self.h))
    def run_query_tags(self):
        _ck(lib().pgi_batch_run_query_tags(self.h))
    def stage_ms(self):
        ms = (ctypes.c_double * 8)(); _ck(lib().pgi_batch_stage_ms(self.h, ctypes.byref(ms)))
        return list(ms)
    def device_result(self):
        res = Result(); _ck(lib().pgi_batch_device_result(self.h, ctypes.byref(res))); return res
    def fetch(self, as_dict=True, text=False):
        res = Result(); _ck(lib().pgi_batch_fetch(self.h, ctypes.byref(res)))
        return _result_to_dict(res, lib(), self.index.has_tags, text=text) if as_dict else res

def random_access_gbs(device=0, buf_bytes=8 << 30, bytes_per_access=32, n_access=1 << 28, chain=8):
    g = ctypes.c_double(); _ck(lib().pgi_random_access_gbs(device, buf_bytes, bytes_per_access, n_access, chain, ctypes.byref(g)))
    return g.value

def l2_fetch_granularity(device=0, set_bytes=0):
    v = ctypes.c_int(); _ck(lib().pgi_l2_fetch_granularity(device, set_bytes, ctypes.byref(v))); return v.value
