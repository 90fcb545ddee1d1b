"""CPU: the drop-in boundary. libpgi.so loads and exports every symbol include/pgi.h declares; on a box without a
CUDA device every compute entry point FAILS LOUDLY (PGI_ENODEVICE) — there is no CPU fallback to fall into.
No compute calls are made here."""
import ctypes, os, re
import numpy as np, pytest
import datasets
from conftest import load_package, ROOT

def header_symbols():
    h = open(os.path.join(ROOT, "include", "pgi.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(pgi_[a-z0-9_]+)\s*\(", h)))

def test_library_exports_every_declared_symbol():
    capi = load_package().capi
    L = capi.lib()
    syms = header_symbols()
    assert len(syms) >= 22
    for s in syms:
        assert hasattr(L, s), "libpgi.so does not export " + s
    assert sorted(capi.SYMBOLS) == syms, "capi.SYMBOLS is out of sync with include/pgi.h"

def test_structs_match_header_layout():
    capi = load_package().capi
    assert ctypes.sizeof(capi.Result) == 12 * 8                 # pgi_result: 3 counts, 5 pointers, extensions, seed_lookups, blocks_fetched, owner
    assert ctypes.sizeof(capi.IndexInfo) == 20 * 8 + 8 * 8 + 256
    assert ctypes.sizeof(capi.LocateResult) == 7 * 8

def test_product_does_not_link_the_oracle():
    """The product library must not contain or load the oracle / hostsim (they are checkers)."""
    capi = load_package().capi
    blob = open(capi.LIB_PATH, "rb").read()
    assert b"orc_find_mems" not in blob and b"sim_find_mems" not in blob and b"liborc" not in blob
    d = os.path.join(ROOT, "pangenome-index_b200", "csrc")
    src = "".join(open(os.path.join(d, f)).read() for f in os.listdir(d) if f.endswith((".cu", ".cuh", ".hpp", ".h")))
    assert "oracle/" not in src and "orc_" not in src

def test_no_cpu_fallback_without_a_device():
    capi = load_package().capi
    L = capi.lib()
    if L.pgi_device_count() > 0:
        pytest.skip("a CUDA device is present; the no-device behaviour is checked on the build box")
    with pytest.raises(capi.PgiError) as e:
        capi.Index(datasets.gold("xy.ri"))
    assert "-3" in str(e.value) and "no usable CUDA device" in str(e.value)
    g = ctypes.c_double()
    assert L.pgi_random_access_gbs(0, 1 << 20, 32, 1 << 10, 1, ctypes.byref(g)) == -3

def test_argument_errors_do_not_crash():
    capi = load_package().capi
    L = capi.lib()
    assert L.pgi_index_create(None, 0, None, 0, 0, 0, None) == -1          # PGI_EINVAL before any device work
    assert L.pgi_batch_run(None, 20, 1) == -1
    assert L.pgi_index_info(None, None) == -1
    L.pgi_index_free(None); L.pgi_batch_free(None); L.pgi_result_free(None); L.pgi_free(None)
    assert b"null" in L.pgi_last_error()

def test_format_text_is_host_side():
    """pgi_format_text (stdout grammar of src/find_mems.cpp:115-139) needs no device: format a hand-made result."""
    capi = load_package().capi; L = capi.lib()
    mem_off = np.array([0, 1, 1], dtype=np.uint64); mems = np.array([[3, 10, 1381, 4]], dtype=np.uint64)
    tag_off = np.array([0, 3], dtype=np.uint64); tags = np.array([62467, 128008, 203779], dtype=np.uint64); nr = np.array([3], dtype=np.uint64)
    r = capi.Result(2, 1, 3, mem_off.ctypes.data, mems.ctypes.data, tag_off.ctypes.data, tags.ctypes.data, nr.ctypes.data, 0, 0, 0, None)
    ln = ctypes.c_uint64(); p = L.pgi_format_text(ctypes.byref(r), 2, 1, ctypes.byref(ln))
    txt = ctypes.string_at(p, ln.value); L.pgi_free(p)
    assert txt == (b"Seq: 2\nMEM START: 3, MEM END: 10 BWT START: 1381 SIZE: 4\nNumber of unique positions: 3\n"
                   b"62467, 128008, 203779, \n\nSeq: 3\n\n")

def test_format_text_matches_a_plain_formatter_on_random_results():
    """pgi_format_text is host code (exact-size pass + in-place pass over read ranges on all host threads): its bytes against
    a plain Python rendering of the same grammar (src/find_mems.cpp:115-139, src/tag_arrays.cpp:885-889) on random results
    with every decimal length (0, 9, 10, 99, ..., 2^64 - 1), negative sizes, reads without MEMs, MEMs without tags; one and
    several threads."""
    import ctypes, os
    import numpy as np
    pkg = load_package(); capi = pkg.capi; L = capi.lib()
    rng = np.random.default_rng(3)
    edge = [0, 9, 10, 99, 100, 999, 1000, 10**9 - 1, 10**9, 2**32 - 1, 2**32, 10**10, 2**40 - 1, 10**18, 10**19 - 1, 10**19, 2**64 - 1]
    def rnd(n):
        v = rng.integers(0, 2**63, n, dtype=np.uint64) >> rng.integers(0, 63, n).astype(np.uint64)
        v[rng.random(n) < 0.2] = rng.choice(np.array(edge, dtype=np.uint64), int((rng.random(n) < 0.2).sum() or 1))[0]
        idx = rng.integers(0, n, min(n, len(edge))); v[idx] = np.array(edge, dtype=np.uint64)[: len(idx)]
        return v
    L.pgi_format_text.restype = ctypes.c_void_p
    L.pgi_format_text.argtypes = [ctypes.POINTER(capi.Result), ctypes.c_uint64, ctypes.c_int, ctypes.POINTER(ctypes.c_uint64)]
    for n_reads, threads in ((37, "1"), (20000, "1"), (20000, "5")):
        cnt = rng.integers(0, 4, n_reads); cnt[rng.random(n_reads) < 0.3] = 0
        mem_off = np.zeros(n_reads + 1, dtype=np.uint64); mem_off[1:] = np.cumsum(cnt); nm = int(mem_off[-1])
        mems = np.zeros((nm, 4), dtype=np.uint64)
        mems[:, 0] = rnd(nm); mems[:, 1] = rnd(nm); mems[:, 2] = rnd(nm); mems[:, 3] = rnd(nm)       # size read as int64: negatives occur
        tcnt = rng.integers(0, 4, nm); tag_off = np.zeros(nm + 1, dtype=np.uint64); tag_off[1:] = np.cumsum(tcnt); nt = int(tag_off[-1])
        tags = rnd(max(nt, 1))[:nt] if nt else np.zeros(0, dtype=np.uint64)
        n_runs = np.zeros(nm, dtype=np.uint64)
        res = capi.Result(); res.n_reads = n_reads; res.n_mems = nm; res.n_tags = nt
        res.mem_off = mem_off.ctypes.data; res.mems = mems.ctypes.data; res.tag_off = tag_off.ctypes.data; res.tags = tags.ctypes.data; res.n_runs = n_runs.ctypes.data
        os.environ["PGI_FORMAT_THREADS"] = threads
        try:
            for with_tags in (1, 0):
                ln = ctypes.c_uint64(); p = L.pgi_format_text(ctypes.byref(res), 1000, with_tags, ctypes.byref(ln))
                assert p
                got = ctypes.string_at(p, ln.value); L.pgi_free(ctypes.c_void_p(p))
                out = []
                for i in range(n_reads):
                    out.append("Seq: %d\n" % (1000 + i))
                    for m in range(int(mem_off[i]), int(mem_off[i + 1])):
                        sz = int(mems[m, 3]); sz = sz - (1 << 64) if sz >= (1 << 63) else sz
                        out.append("MEM START: %d, MEM END: %d BWT START: %d SIZE: %d\n" % (int(mems[m, 0]), int(mems[m, 1]), int(mems[m, 2]), sz))
                        if with_tags:
                            ts = tags[int(tag_off[m]):int(tag_off[m + 1])]
                            out.append("Number of unique positions: %d\n" % len(ts)); out.append("".join("%d, " % int(t) for t in ts) + "\n")
                    out.append("\n")
                assert got == "".join(out).encode(), (n_reads, threads, with_tags)
        finally:
            os.environ.pop("PGI_FORMAT_THREADS", None)
