"""ctypes binding of the CPU oracle (oracle/liborc.so) — test infrastructure; the checker, never the product."""
import ctypes, os, subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_DIR = os.path.join(ROOT, "oracle")
_lib = None
_ref = None
REF_SO = os.path.join(_DIR, "_ref", "libpgiref.so")

class OrcResult(ctypes.Structure):
    _fields_ = [("n_reads", ctypes.c_uint64), ("n_mems", ctypes.c_uint64), ("n_tags", ctypes.c_uint64),
                ("mem_off", ctypes.POINTER(ctypes.c_uint64)), ("mems", ctypes.POINTER(ctypes.c_uint64)),
                ("tag_off", ctypes.POINTER(ctypes.c_uint64)), ("tags", ctypes.POINTER(ctypes.c_uint64)),
                ("n_runs", ctypes.POINTER(ctypes.c_uint64)), ("counters", ctypes.c_uint64 * 8)]

def ref_available():
    """oracle/_ref/libpgiref.so = the reference's own r-index.cpp / tag_arrays.cpp / find_all_mems compiled against
    oracle/ref_shim (oracle/Makefile). Built where /root/reference is mounted; travels prebuilt to the GPU box."""
    if not os.path.exists(REF_SO) and os.path.exists("/root/reference/src/r-index.cpp"):
        subprocess.call(["make", "-s", "-C", _DIR, "_ref/libpgiref.so"])
    return os.path.exists(REF_SO)

def ref_lib():
    global _ref
    if _ref is None:
        if not ref_available():
            raise RuntimeError("oracle/_ref/libpgiref.so is not built (needs /root/reference)")
        _ref = _bind(ctypes.CDLL(REF_SO), ref=True)
    return _ref

def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_DIR, "liborc.so")
        if not os.path.exists(path):
            subprocess.check_call(["make", "-s", "-C", _DIR, "liborc.so"])
        _lib = _bind(ctypes.CDLL(path), ref=False)
    return _lib

def _bind(L, ref):
    if True:
        L.orc_last_error.restype = ctypes.c_char_p
        L.orc_index_load.restype = ctypes.c_void_p
        L.orc_index_load.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_int,
                                     ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
        L.orc_index_free.argtypes = [ctypes.c_void_p]
        L.orc_index_info.argtypes = [ctypes.c_void_p] * 4
        L.orc_extend_batch.argtypes = [ctypes.c_void_p, ctypes.c_uint64] + [ctypes.c_void_p] * 4
        L.orc_rank_batch.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p]
        L.orc_tag_query.restype = ctypes.c_uint64
        L.orc_tag_query.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint64,
                                    ctypes.POINTER(ctypes.c_uint64)]
        L.orc_find_mems.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64,
                                    ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(OrcResult)]
        if not ref:
            L.orc_query_tags.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int, ctypes.c_int,
                                         ctypes.POINTER(OrcResult)]
            L.orc_format_query_tags.restype = ctypes.c_void_p
            L.orc_format_query_tags.argtypes = [ctypes.POINTER(OrcResult), ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64)]
            L.orc_format_text.restype = ctypes.c_void_p
            L.orc_format_text.argtypes = [ctypes.POINTER(OrcResult), ctypes.c_uint64, ctypes.c_int, ctypes.POINTER(ctypes.c_uint64)]
        else:
            L.ref_time_find_mems.restype = ctypes.c_double
            L.ref_time_find_mems.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64,
                                             ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_uint64)]
            L.ref_find_mems_text.restype = ctypes.c_void_p
            L.ref_find_mems_text.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int,
                                             ctypes.POINTER(ctypes.c_uint64)]
            L.ref_query_tags_text.restype = ctypes.c_void_p
            L.ref_query_tags_text.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64)]
            L.ref_build_rindex.restype = ctypes.c_void_p
            L.ref_build_rindex.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.POINTER(ctypes.c_uint64)]
        L.orc_locate_batch.restype = ctypes.c_uint64
        L.orc_locate_batch.argtypes = [ctypes.c_void_p, ctypes.c_uint64] + [ctypes.c_void_p] * 4 + [ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint64]
        L.orc_decompress_sa.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.orc_result_free.argtypes = [ctypes.POINTER(OrcResult)]
        L.orc_free.argtypes = [ctypes.c_void_p]
    return L

def _np(ptr, n, dtype=np.uint64):
    if n == 0:
        return np.zeros(0, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)

class Index:
    """Oracle index: FastLocate (+ TagArray). tag_layout 0 = sdsl-compact, 1 = ByteCode+len9."""
    def __init__(self, ri: bytes, tags: bytes = None, tag_layout=0, impl="port"):
        """impl "port": oracle/liborc.so (the restatement). impl "ref": oracle/_ref/libpgiref.so (the reference's own code)."""
        self.impl = impl; self.L = L = ref_lib() if impl == "ref" else lib(); a = ctypes.c_uint64(); b = ctypes.c_uint64()
        self.tag_layout = tag_layout
        self.h = L.orc_index_load(ri, len(ri), tags, len(tags) if tags else 0, tag_layout, ctypes.byref(a), ctypes.byref(b))
        if not self.h:
            raise RuntimeError(L.orc_last_error().decode())
        self.ri_consumed, self.tags_consumed = a.value, b.value
        self.has_tags = bool(tags)
        info = (ctypes.c_uint64 * 16)(); C = (ctypes.c_uint64 * 8)(); sm = (ctypes.c_uint8 * 256)()
        L.orc_index_info(self.h, info, C, sm)
        self.info = list(info); self.n = info[0]; self.sigma = info[1]
        self.C = list(C)[: self.sigma]; self.sym_map = bytes(sm)
    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.L.orc_index_free(self.h); self.h = None
        except Exception:
            pass
    def extend_batch(self, iv3, sym, direction):
        iv3 = np.ascontiguousarray(iv3, dtype=np.uint64); sym = np.ascontiguousarray(sym, dtype=np.uint8)
        direction = np.ascontiguousarray(direction, dtype=np.uint8); out = np.zeros_like(iv3)
        self.L.orc_extend_batch(self.h, len(sym), iv3.ctypes.data, sym.ctypes.data, direction.ctypes.data, out.ctypes.data)
        return out
    def rank_batch(self, pos):
        pos = np.ascontiguousarray(pos, dtype=np.uint64); out = np.zeros((len(pos), 8), dtype=np.uint64)
        self.L.orc_rank_batch(self.h, len(pos), pos.ctypes.data, out.ctypes.data)
        return out
    def search(self, pattern: bytes):
        iv = np.array([[0, 0, self.n]], dtype=np.uint64)
        for ch in reversed(pattern):
            iv = self.extend_batch(iv, [ch], [0])
        return tuple(int(x) for x in iv[0])
    def tag_query(self, start, end, cap=1 << 20):
        out = np.zeros(cap, dtype=np.uint64); n = ctypes.c_uint64()
        nr = self.L.orc_tag_query(self.h, start, end, out.ctypes.data, cap, ctypes.byref(n))
        return nr, out[: n.value].copy()
    def find_mems(self, reads, read_off, min_len, min_occ, with_tags=None, threads=0, count_work=False, text=True):
        """reads: uint8 array (concatenated), read_off: uint64 [n+1] -> dict of numpy arrays."""
        reads = np.ascontiguousarray(reads, dtype=np.uint8); read_off = np.ascontiguousarray(read_off, dtype=np.uint64)
        if with_tags is None:
            with_tags = self.has_tags
        import time
        res = OrcResult(); t0 = time.perf_counter()
        if self.L.orc_find_mems(self.h, reads.ctypes.data, read_off.ctypes.data, len(read_off) - 1, min_len, min_occ,
                               int(with_tags), threads, int(count_work), ctypes.byref(res)):
            raise RuntimeError(self.L.orc_last_error().decode())
        seconds = time.perf_counter() - t0          # the C call only: MEMs + tags into binary arrays, no text
        out = dict(seconds=seconds, n_reads=res.n_reads, n_mems=res.n_mems, n_tags=res.n_tags,
                   mem_off=_np(res.mem_off, res.n_reads + 1), mems=_np(res.mems, 4 * res.n_mems).reshape(-1, 4),
                   tag_off=_np(res.tag_off, res.n_mems + 1), tags=_np(res.tags, res.n_tags), n_runs=_np(res.n_runs, res.n_mems),
                   counters=dict(zip(["E", "D", "M", "tag_runs", "tag_skipped", "tag_item_sectors"], list(res.counters)[:6])))
        if text and self.impl != "ref":
            ln = ctypes.c_uint64()
            p = self.L.orc_format_text(ctypes.byref(res), 1, int(with_tags), ctypes.byref(ln))
            out["text"] = ctypes.string_at(p, ln.value); self.L.orc_free(p)
        self.L.orc_result_free(ctypes.byref(res))
        return out
    # ---- impl "ref" only
    def ref_text(self, reads_text: bytes, min_len, min_occ, verbatim=True):
        """stdout of the reference's own find_mems main loop (src/find_mems.cpp:93-139) over newline-separated reads."""
        ln = ctypes.c_uint64()
        p = self.L.ref_find_mems_text(self.h, reads_text, len(reads_text), min_len, min_occ, int(verbatim), ctypes.byref(ln))
        if not p: raise RuntimeError(self.L.orc_last_error().decode())
        t = ctypes.string_at(p, ln.value); self.L.orc_free(p); return t
    def ref_query_tags_text(self, reads_text: bytes):
        ln = ctypes.c_uint64()
        p = self.L.ref_query_tags_text(self.h, reads_text, len(reads_text), ctypes.byref(ln))
        if not p: raise RuntimeError(self.L.orc_last_error().decode())
        t = ctypes.string_at(p, ln.value); self.L.orc_free(p); return t
    def ref_time(self, reads, read_off, min_len, min_occ, with_tags=True, threads=0):
        """Seconds of the reference's per-read work (find_all_mems + tag query of every MEM), OpenMP over reads. -> (s, n_mems)"""
        reads = np.ascontiguousarray(reads, dtype=np.uint8); read_off = np.ascontiguousarray(read_off, dtype=np.uint64); nm = ctypes.c_uint64()
        dt = self.L.ref_time_find_mems(self.h, reads.ctypes.data, read_off.ctypes.data, len(read_off) - 1, min_len, min_occ, int(with_tags and self.has_tags), threads, ctypes.byref(nm))
        return dt, nm.value

def ref_build_rindex(rlbwt_path, encoded=True):
    """The reference's FastLocate(rlbwt) constructor + serialize[_encoded] -> .ri bytes (src/build_rindex.cpp)."""
    L = ref_lib(); ln = ctypes.c_uint64()
    p = L.ref_build_rindex(rlbwt_path.encode(), int(encoded), ctypes.byref(ln))
    if not p: raise RuntimeError(L.orc_last_error().decode())
    b = ctypes.string_at(p, ln.value); L.orc_free(p); return b

class _Unused:
    pass

def _query_tags(self, reads, read_off, with_tags=None, threads=0, text=True):
    """query_tags path (src/query_tags.cpp:87-109): one record per read {0, len, range.first, size}."""
    reads = np.ascontiguousarray(reads, dtype=np.uint8); read_off = np.ascontiguousarray(read_off, dtype=np.uint64)
    if with_tags is None:
        with_tags = self.has_tags
    import time
    res = OrcResult(); t0 = time.perf_counter()
    if self.L.orc_query_tags(self.h, reads.ctypes.data, read_off.ctypes.data, len(read_off) - 1, int(with_tags), threads, ctypes.byref(res)):
        raise RuntimeError(self.L.orc_last_error().decode())
    seconds = time.perf_counter() - t0
    out = dict(seconds=seconds, n_reads=res.n_reads, n_mems=res.n_mems, n_tags=res.n_tags,
               mem_off=_np(res.mem_off, res.n_reads + 1), mems=_np(res.mems, 4 * res.n_mems).reshape(-1, 4),
               tag_off=_np(res.tag_off, res.n_mems + 1), tags=_np(res.tags, res.n_tags), n_runs=_np(res.n_runs, res.n_mems))
    if text:
        ln = ctypes.c_uint64()
        p = self.L.orc_format_query_tags(ctypes.byref(res), 0, ctypes.byref(ln))
        out["text"] = ctypes.string_at(p, ln.value); self.L.orc_free(p)
    self.L.orc_result_free(ctypes.byref(res))
    return out
Index.query_tags = _query_tags

def _locate_batch(self, first, second, want_sa=False):
    """locate_encoded (r-index.cpp:1300-1343) per range -> (offsets [n+1], sorted unique sequence ids[, SA values row by row])."""
    first = np.ascontiguousarray(first, dtype=np.uint64); second = np.ascontiguousarray(second, dtype=np.uint64)
    n = len(first); sizes = np.where(second >= first, second - first + np.uint64(1), 0).astype(np.uint64); tot = int(sizes.sum())
    off = np.zeros(n + 1, dtype=np.uint64); out = np.zeros(max(1, tot), dtype=np.uint64); sa = np.zeros(max(1, tot), dtype=np.uint64)
    k = self.L.orc_locate_batch(self.h, n, first.ctypes.data, second.ctypes.data, off.ctypes.data, out.ctypes.data, tot,
                               sa.ctypes.data if want_sa else None, tot)
    return (off, out[:k].copy(), sa[:tot].copy()) if want_sa else (off, out[:k].copy())
def _decompress_sa(self):
    out = np.zeros(self.n, dtype=np.uint64); self.L.orc_decompress_sa(self.h, out.ctypes.data); return out
Index.locate_batch = _locate_batch
Index.decompress_sa = _decompress_sa

def pack_reads(reads):
    """list of bytes -> (uint8 concatenation, uint64 offsets)."""
    off = np.zeros(len(reads) + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(r) for r in reads], dtype=np.uint64)
    buf = np.frombuffer(b"".join(reads), dtype=np.uint8) if len(reads) else np.zeros(0, dtype=np.uint8)
    return buf.copy(), off

def algorithmic_bytes(counters):
    """128*D + 96*M + 32*sum ceil(8*(skip+runs)/32)."""
    return 128 * counters["D"] + 96 * counters["M"] + 32 * counters["tag_item_sectors"]
