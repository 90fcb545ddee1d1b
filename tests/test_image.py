"""Cached flattened index image (the GPU counterpart of the reference's MemoryMappedIndexManager /
ChunkedMemoryMappedIndexManager, src/find_mems_memory_mapped.cpp:36-45, src/find_mems_chunked.cpp:29-50).
CPU: pgi_image_write / pgi_image_info are host-only (flatten + file layout + validation). GPU: an index opened from an
image, with any chunk size, answers every query exactly like one created from the SDSL files (and like the oracle)."""
import ctypes, os
import numpy as np, pytest
import orc, datasets
from conftest import load_package

def _cases():
    s = datasets.synth_index()
    return {"xy_bytecode": (datasets.gold("xy.ri"), datasets.gold("xy_bidirectional_compressed.tags"), 1, datasets.gold("contigs_xy")),
            "synth_compact": (s["ri"], s["tags0"], 0, s["text"]),
            "ri_only": (datasets.gold("xy.ri"), None, 0, datasets.gold("contigs_xy"))}

def test_image_write_and_info_are_host_only(tmp_path):
    capi = load_package().capi
    for name, (ri, tg, layout, _) in _cases().items():
        p = str(tmp_path / (name + ".pgimg"))
        capi.image_write(ri, tg, layout, p)
        f = capi.image_info(p, verify=True)
        assert f["file_bytes"] == os.path.getsize(p) and f["file_bytes"] % 4096 == 0
        assert f["has_ri"] == 1 and f["has_tags"] == (1 if tg else 0) and f["tag_layout"] == (layout if tg else 0)
        assert f["src_ri_bytes"] == len(ri) and f["src_ri_sum"] == capi.image_source_sum(ri)
        assert f["src_tag_bytes"] == (len(tg) if tg else 0)
        assert f["n_dev_blocks"] * 64 <= f["payload_bytes"] < f["file_bytes"]
        if name != "synth_compact":
            assert f["n"] == 8022 and f["sigma"] == 5 and f["is_encoded"] == 0        # SURVEY App. D facts of xy.ri
        p2 = str(tmp_path / (name + ".again"))
        capi.image_write(ri, tg, layout, p2)
        assert open(p, "rb").read() == open(p2, "rb").read()                         # deterministic
        assert not os.path.exists(p + ".tmp")

def test_image_rejects_corruption(tmp_path):
    capi = load_package().capi; L = capi.lib()
    ri, tg, layout, _ = _cases()["xy_bytecode"]
    p = str(tmp_path / "a.pgimg"); capi.image_write(ri, tg, layout, p)
    blob = bytearray(open(p, "rb").read()); f = capi.ImageInfo()
    def rc_of(b, verify=0):
        q = str(tmp_path / "bad.pgimg"); open(q, "wb").write(bytes(b))
        return L.pgi_image_info(os.fsencode(q), verify, ctypes.byref(f))
    assert rc_of(blob, 1) == 0
    assert rc_of(blob[:100]) == -2 and b"too short" in L.pgi_last_error()
    assert rc_of(blob[:-4096]) == -2 and b"truncated" in L.pgi_last_error()
    b = bytearray(blob); b[0] ^= 1
    assert rc_of(b) == -2 and b"magic" in L.pgi_last_error()
    b = bytearray(blob); b[200] ^= 0x40                                             # a header scalar
    assert rc_of(b) == -2 and b"checksum" in L.pgi_last_error()
    b = bytearray(blob); b[8192 + 17] ^= 0x08                                       # payload: caught only when verifying
    assert rc_of(b, 0) == 0
    assert rc_of(b, 1) == -2 and b"section checksum" in L.pgi_last_error()
    assert L.pgi_image_info(os.fsencode(str(tmp_path / "missing")), 0, ctypes.byref(f)) == -2
    assert L.pgi_image_write(None, 0, None, 0, 0, b"/tmp/x") == -1
    assert L.pgi_image_write(ri, len(ri), tg, len(tg), 0, os.fsencode(p)) == -2    # wrong tag layout: FormatError, file untouched
    assert open(p, "rb").read() == bytes(blob)
    if L.pgi_device_count() == 0:
        h = ctypes.c_void_p()
        assert L.pgi_index_open_image(os.fsencode(p), 0, 0, 0, ctypes.byref(h), None) == -3   # no CPU path behind an image either

@pytest.mark.gpu
@pytest.mark.parametrize("case", ["xy_bytecode", "synth_compact", "ri_only"])
def test_index_from_image_equals_index_from_files(tmp_path, case):
    capi = load_package().capi
    ri, tg, layout, text = _cases()[case]
    p = str(tmp_path / "x.pgimg"); capi.image_write(ri, tg, layout, p)
    a = capi.Index(ri, tg, layout)
    reads, off = datasets.sample_reads(text, 3000, 150, 0.02, seed=11)
    want = a.find_mems(reads, off, 20, 1, text=True)
    o = orc.Index(ri, tg, layout).find_mems(reads, off, 20, 1)
    assert want["text"] == o["text"]
    file_bytes = os.path.getsize(p)
    for chunk, verify in [(0, False), (4096, True), (12288, False), (1 << 30, False)]:
        b = capi.Index.open_image(p, chunk_bytes=chunk, verify=verify)
        assert b.info == a.info
        assert b.image["chunk_bytes"] == (chunk or 64 << 20)
        if chunk == 4096: assert b.image["chunks_copied"] >= b.image["payload_bytes"] // 4096
        if chunk == 1 << 30: assert b.image["chunks_copied"] <= 14 and b.image["file_bytes"] == file_bytes
        got = b.find_mems(reads, off, 20, 1, text=True)
        for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"):
            assert np.array_equal(got[k], want[k]), (chunk, k)
        assert got["text"] == want["text"] and got["extensions"] == want["extensions"]
        if case == "synth_compact":
            qa = a.query_tags(reads, off); qb = b.query_tags(reads, off)
            assert qa["text"] == qb["text"]
        if a.info["has_locate"]:
            m = want["mems"][:200]
            la = a.locate_batch(m[:, 2], m[:, 2] + m[:, 3] - 1); lb = b.locate_batch(m[:, 2], m[:, 2] + m[:, 3] - 1)
            assert all(np.array_equal(x, y) for x, y in zip(la, lb))
        b.close()
    a.close()

@pytest.mark.gpu
def test_out_of_core_run_blocks_in_host_memory(monkeypatch):
    """SURVEY.md section 8 f4 (second half): with the run-block array kept in pinned host memory and read over PCIe
    (PGI_HOST_BLOCKS=1 forces what an index larger than the free HBM triggers by itself) every query answers the same."""
    import orc, datasets
    capi = load_package().capi
    s = datasets.synth_index(ref_len=20000, n_hap=8, seed=21, encoded=True)
    reads, off = datasets.sample_reads(s["text"], 4000, 150, 0.01, seed=3)
    R = orc.Index(s["ri"], s["tags0"], 0).find_mems(reads, off, 20, 1)
    monkeypatch.setenv("PGI_HOST_BLOCKS", "1")
    g = capi.Index(s["ri"], s["tags0"], 0)
    assert g.info["host_resident_bytes"] == g.info["n_dev_blocks"] * 64
    G = g.find_mems(reads, off, 20, 1)
    assert G["text"] == R["text"] and np.array_equal(G["mems"], R["mems"]) and np.array_equal(G["tags"], R["tags"])
    with pytest.raises(capi.PgiError): g.clone_to(0)
    monkeypatch.delenv("PGI_HOST_BLOCKS")
    assert capi.Index(s["ri"], s["tags0"], 0).info["host_resident_bytes"] == 0
