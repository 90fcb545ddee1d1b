"""ctypes binding of the CPU data-prep tools (prep/libpgi_prep.so): multi-string BWT -> .rl_bwt,
r-index construction + serialisation (legacy / encoded), tag files from per-row graph positions, read
sampling. These surround the hot path; they are not on it."""
import ctypes, os, subprocess
import numpy as np

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "prep")
_lib = None

def build():
    subprocess.check_call(["make", "-s", "-C", _DIR])

def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_DIR, "libpgi_prep.so")
        if not os.path.exists(path):
            build()
        L = ctypes.CDLL(path)
        L.prep_last_error.restype = ctypes.c_char_p
        _lib = L
    return _lib

def _take(L, p, n):
    out = ctypes.string_at(p, n.value)
    L.prep_free(p)
    return out

def _err(L):
    raise RuntimeError(L.prep_last_error().decode())

def bwt_from_text(text: bytes) -> bytes:
    """Newline-terminated sequences -> .rl_bwt file bytes (grlBWT layout)."""
    L = lib(); p = ctypes.c_void_p(); n = ctypes.c_uint64()
    if L.prep_bwt_from_text(text, ctypes.c_uint64(len(text)), ctypes.byref(p), ctypes.byref(n), None): _err(L)
    return _take(L, p, n)

def rindex_from_rlbwt(rlbwt: bytes, encoded=True, with_samples=False, want_rows=False):
    """-> (ri_bytes, info dict[, row_textpos uint64 array])."""
    L = lib(); p = ctypes.c_void_p(); n = ctypes.c_uint64(); info = (ctypes.c_uint64 * 8)()
    rows = None
    if want_rows:
        # n is not known before parsing; compute it from the records
        sb, fb = np.frombuffer(rlbwt[:16], dtype=np.uint64)
        rec = np.frombuffer(rlbwt[16:], dtype=np.uint8).reshape(-1, int(sb + fb))
        lens = np.zeros(len(rec), dtype=np.uint64)
        for i in range(int(fb)):
            lens |= rec[:, int(sb) + i].astype(np.uint64) << np.uint64(8 * i)
        rows = np.zeros(int(lens.sum()), dtype=np.uint64)
    rp = rows.ctypes.data_as(ctypes.c_void_p) if rows is not None else None
    if L.prep_rindex_from_rlbwt(rlbwt, ctypes.c_uint64(len(rlbwt)), int(encoded), int(with_samples),
                                ctypes.byref(p), ctypes.byref(n), rp, info): _err(L)
    d = dict(n=info[0], n_seq=info[1], max_length=info[2], total_runs=info[3], sigma=info[4])
    ri = _take(L, p, n)
    return (ri, d, rows) if want_rows else (ri, d)

def tags_from_rows(row_value: np.ndarray, layout=0):
    """Per-BWT-row encoded graph positions -> (tags file bytes, number of tag runs)."""
    L = lib(); p = ctypes.c_void_p(); n = ctypes.c_uint64(); nr = ctypes.c_uint64()
    rv = np.ascontiguousarray(row_value, dtype=np.uint64)
    if L.prep_tags_from_rows(rv.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint64(len(rv)), int(layout),
                             ctypes.byref(p), ctypes.byref(n), ctypes.byref(nr)): _err(L)
    return _take(L, p, n), nr.value

def sample_reads(text: bytes, n_reads: int, read_len: int, sub=0.0, indel=0.0, seed=1) -> np.ndarray:
    """-> uint8 array [n_reads, read_len]."""
    L = lib(); out = np.zeros((n_reads, read_len), dtype=np.uint8)
    L.prep_sample_reads.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64,
                                    ctypes.c_double, ctypes.c_double, ctypes.c_uint64, ctypes.c_void_p]
    if L.prep_sample_reads(text, len(text), n_reads, read_len, float(sub), float(indel), seed, out.ctypes.data): _err(L)
    return out

def tags_from_runs(value: np.ndarray, length: np.ndarray, layout=0):
    """(value, length) tag runs -> (tags file bytes, number of runs after splitting at 511)."""
    L = lib(); p = ctypes.c_void_p(); n = ctypes.c_uint64(); nr = ctypes.c_uint64()
    v = np.ascontiguousarray(value, dtype=np.uint64); l = np.ascontiguousarray(length, dtype=np.uint64)
    if L.prep_tags_from_runs(v.ctypes.data_as(ctypes.c_void_p), l.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint64(len(v)), int(layout),
                             ctypes.byref(p), ctypes.byref(n), ctypes.byref(nr)): _err(L)
    return _take(L, p, n), nr.value

def rlbwt_from_runs(sym: np.ndarray, length: np.ndarray) -> bytes:
    """(symbol byte, length) BWT runs -> .rl_bwt file bytes (deps/grlBWT/include/bwt_io.h:40-58)."""
    length = np.ascontiguousarray(length, dtype=np.uint64); sym = np.ascontiguousarray(sym, dtype=np.uint8)
    mx = int(length.max()) if len(length) else 1
    fb = max(1, (mx.bit_length() + 7) // 8)
    rec = np.zeros((len(sym), 1 + fb), dtype=np.uint8)
    rec[:, 0] = sym
    for i in range(fb):
        rec[:, 1 + i] = ((length >> np.uint64(8 * i)) & np.uint64(0xFF)).astype(np.uint8)
    return np.array([1, fb], dtype=np.uint64).tobytes() + rec.tobytes()
