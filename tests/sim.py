"""ctypes binding of tests/hostsim (product device arithmetic compiled as plain C++; test-only)."""
import ctypes, os, subprocess
import numpy as np

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hostsim")
_lib = None

def lib():
    global _lib
    if _lib is None:
        subprocess.check_call(["make", "-s", "-C", _DIR])
        L = ctypes.CDLL(os.path.join(_DIR, "libpgi_hostsim.so"))
        L.sim_last_error.restype = ctypes.c_char_p
        L.sim_create.restype = ctypes.c_void_p
        L.sim_create.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_int]
        L.sim_free.argtypes = [ctypes.c_void_p]
        L.sim_info.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.sim_extend_batch.argtypes = [ctypes.c_void_p, ctypes.c_uint64] + [ctypes.c_void_p] * 4
        L.sim_find_mems.restype = ctypes.c_uint64
        L.sim_find_mems.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64,
                                    ctypes.c_void_p, ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64)]
        L.sim_tag_query.restype = ctypes.c_uint64
        L.sim_tag_query.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint64,
                                    ctypes.POINTER(ctypes.c_uint64)]
        L.sim_count.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_void_p]
        L.sim_locate.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_void_p]
        _lib = L
    return _lib

class Index:
    def __init__(self, ri, tags=None, tag_layout=0):
        L = lib()
        self.h = L.sim_create(ri, len(ri), tags, len(tags) if tags else 0, tag_layout)
        if not self.h:
            raise RuntimeError(L.sim_last_error().decode())
        info = (ctypes.c_uint64 * 16)(); L.sim_info(self.h, info)
        keys = ["n", "sigma", "n_blocks", "bshift", "has_quirk", "n_file_runs", "n_ref_blocks", "encoded",
                "tag_m", "tag_shift", "tag_low_bytes", "starts_regular", "n_crowded", "n_buckets"]
        self.info = dict(zip(keys, list(info)))
    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().sim_free(self.h); self.h = None
        except Exception:
            pass
    def extend_batch(self, iv3, sym, direction):
        iv3 = np.ascontiguousarray(iv3, dtype=np.uint64); sym = np.ascontiguousarray(sym, dtype=np.uint8)
        direction = np.ascontiguousarray(direction, dtype=np.uint8); out = np.zeros_like(iv3)
        lib().sim_extend_batch(self.h, len(sym), iv3.ctypes.data, sym.ctypes.data, direction.ctypes.data, out.ctypes.data)
        return out
    def set_seed_k(self, k):
        """Build (k > 0) or drop (k = 0) the seed table; returns the K in effect."""
        lib().sim_set_seed_k.restype = ctypes.c_uint32; lib().sim_set_seed_k.argtypes = [ctypes.c_void_p, ctypes.c_uint32]
        return lib().sim_set_seed_k(self.h, k)
    def find_mems(self, read: bytes, min_len, min_occ, cap=4096):
        out = np.zeros((cap, 4), dtype=np.uint64); ne = ctypes.c_uint64()
        n = lib().sim_find_mems(self.h, read, len(read), min_len, min_occ, out.ctypes.data, cap, ctypes.byref(ne))
        return out[:n].copy(), ne.value
    def count(self, read: bytes):
        """-> (range.first, size), (1, 0) without a match, or None when the index refuses counting."""
        out = np.zeros(2, dtype=np.uint64)
        if lib().sim_count(self.h, read, len(read), out.ctypes.data):
            return None
        return int(out[0]), int(out[1])
    def locate(self, first, second):
        """SA values of BWT rows [first, second], or None when the index refuses locate."""
        out = np.zeros(max(1, second - first + 1), dtype=np.uint64)
        if lib().sim_locate(self.h, first, second, out.ctypes.data):
            return None
        return out[: max(0, second - first + 1)]
    def tag_query(self, start, end, cap=1 << 20):
        out = np.zeros(cap, dtype=np.uint64); n = ctypes.c_uint64()
        nr = lib().sim_tag_query(self.h, start, end, out.ctypes.data, cap, ctypes.byref(n))
        return nr, out[: n.value].copy()
