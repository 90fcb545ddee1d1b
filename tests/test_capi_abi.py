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
