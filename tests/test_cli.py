"""The find_mems command line (host/find_mems.cpp: batched multi-GPU driver; host/find_mems_serial.cpp: the reference's
per-read loop written against the facade header). stdout must equal the oracle's rendering of src/find_mems.cpp:115-139,
byte for byte, minus the two trailing timing lines (whose values are run dependent in the reference too)."""
import os, subprocess, sys
import numpy as np, pytest
import orc, datasets
from conftest import ROOT

HOST = os.path.join(ROOT, "pangenome-index_b200", "host")
CLI = os.path.join(HOST, "find_mems"); SERIAL = os.path.join(HOST, "find_mems_serial")

def strip_timing(out: bytes) -> bytes:
    i = out.rfind(b"\nTotal time for finding all MEMs:")
    assert i >= 0 and out.count(b"Total time for") == 2
    return out[:i]

def test_cli_fails_loudly_without_gpu(tmp_path):
    from conftest import load_package
    if load_package().capi.lib().pgi_device_count() > 0:
        pytest.skip("a CUDA device is present")
    p = subprocess.run([CLI, os.path.join(datasets.GOLD, "xy.ri"), os.path.join(datasets.GOLD, "xy_bidirectional_compressed.tags"),
                        os.path.join(datasets.GOLD, "reads.txt"), "5", "1"], capture_output=True)
    assert p.returncode != 0 and b"no CUDA device" in p.stderr and p.stdout == b""
    p = subprocess.run([CLI], capture_output=True)
    assert p.returncode != 0 and b"usage" in p.stderr

@pytest.mark.gpu
@pytest.mark.parametrize("case", ["xy_bytecode", "synth_compact"])
def test_cli_stdout_equals_oracle(tmp_path, case):
    if case == "xy_bytecode":
        ri, tg, layout, text = datasets.gold("xy.ri"), datasets.gold("xy_bidirectional_compressed.tags"), 1, datasets.gold("contigs_xy")
    else:
        s = datasets.synth_index(); ri, tg, layout, text = s["ri"], s["tags0"], 0, s["text"]
    rs = datasets.ragged_reads(text, seed=9, n=300)
    rs = [r for r in rs if b"\n" not in r and b"\0" not in r]          # one read per text line
    reads, off = datasets.sample_reads(text, 500, 150, 0.02, seed=3)
    rs += [reads[int(off[i]):int(off[i + 1])].tobytes() for i in range(500)]
    (tmp_path / "x.ri").write_bytes(ri); (tmp_path / "x.tags").write_bytes(tg)
    body = b"\n".join(rs) + b"\n"
    body = b"\n\n" + body.replace(b"\n", b"\n\n", 5)                   # empty lines are skipped and not numbered
    (tmp_path / "reads.txt").write_bytes(body)
    nonempty = [r for r in rs if r]
    buf, o = orc.pack_reads(nonempty)
    for ml, mo in [(20, 1), (5, 2)]:
        want = orc.Index(ri, tg, layout).find_mems(buf, o, ml, mo)["text"]
        for extra in (["--batch", "97"], ["--batch", "100000"], ["--tags-layout", "bytecode" if layout else "compact"]):
            p = subprocess.run([CLI, str(tmp_path / "x.ri"), str(tmp_path / "x.tags"), str(tmp_path / "reads.txt"), str(ml), str(mo)] + extra, capture_output=True)
            assert p.returncode == 0, p.stderr.decode()
            assert strip_timing(p.stdout) == want, extra
    small = b"\n".join(nonempty[:60]) + b"\n"
    (tmp_path / "small.txt").write_bytes(small)
    sb, so = orc.pack_reads(nonempty[:60])
    want = orc.Index(ri, tg, layout).find_mems(sb, so, 20, 1)["text"]
    p = subprocess.run([SERIAL, str(tmp_path / "x.ri"), str(tmp_path / "x.tags"), str(tmp_path / "small.txt"), "20", "1", "bytecode" if layout else "compact"], capture_output=True)
    assert p.returncode == 0, p.stderr.decode()
    assert p.stdout == want                                            # the facade loop prints no timing lines

@pytest.mark.gpu
def test_cli_rejects_wrong_tag_layout(tmp_path):
    s = datasets.synth_index()
    (tmp_path / "x.ri").write_bytes(s["ri"]); (tmp_path / "x.tags").write_bytes(s["tags0"]); (tmp_path / "r.txt").write_bytes(b"ACGT\n")
    p = subprocess.run([CLI, str(tmp_path / "x.ri"), str(tmp_path / "x.tags"), str(tmp_path / "r.txt"), "2", "1", "--tags-layout", "bytecode"], capture_output=True)
    assert p.returncode != 0 and b"cannot load the index" in p.stderr

@pytest.mark.gpu
def test_query_tags_cli_stdout_equals_oracle(tmp_path):
    """host/query_tags (the same driver built with -DPGI_QUERY_TAGS): stdout of src/query_tags.cpp:87-109, read_index
    counting non-empty lines from 0; reads without a match print nothing on stdout."""
    QT = os.path.join(HOST, "query_tags")
    s = datasets.synth_index(); ri, tg, text = s["ri"], s["tags0"], s["text"]
    rng = np.random.default_rng(4); lines = [l for l in text.split(b"\n") if len(l) > 100]; rs = []
    for _ in range(700):
        L = lines[int(rng.integers(len(lines)))]; ln = int(rng.integers(1, 90)); st = int(rng.integers(0, len(L) - ln))
        r = bytearray(L[st:st + ln])
        if rng.random() < 0.3: r[int(rng.integers(len(r)))] = b"ACGTNx"[int(rng.integers(6))]
        rs.append(bytes(r))
    (tmp_path / "x.ri").write_bytes(ri); (tmp_path / "x.tags").write_bytes(tg)
    (tmp_path / "reads.txt").write_bytes(b"\n" + b"\n".join(rs[:5]) + b"\n\n\n" + b"\n".join(rs[5:]) + b"\n")
    buf, o = orc.pack_reads(rs)
    want = orc.Index(ri, tg, 0).query_tags(buf, o)["text"]
    assert want.count(b"read_index=") > 300
    for extra in ([], ["--batch", "53"]):
        p = subprocess.run([QT, str(tmp_path / "x.ri"), str(tmp_path / "x.tags"), str(tmp_path / "reads.txt")] + extra, capture_output=True)
        assert p.returncode == 0, p.stderr.decode()
        assert p.stdout == want, extra
    (tmp_path / "empty.txt").write_bytes(b"\n\n")
    p = subprocess.run([QT, str(tmp_path / "x.ri"), str(tmp_path / "x.tags"), str(tmp_path / "empty.txt")], capture_output=True)
    assert p.returncode == 1 and b"No reads found in file." in p.stderr

@pytest.mark.gpu
@pytest.mark.parametrize("case", ["xy_bytecode", "synth_compact"])
def test_mmap_and_chunked_cli_stdout_equals_oracle(tmp_path, case):
    """host/find_mems_memory_mapped and host/find_mems_chunked (src/find_mems_memory_mapped.cpp, src/find_mems_chunked.cpp):
    the per-read text of find_mems, then the reference's trailer lines. The first run writes the cached device image,
    later runs map it; a changed source file invalidates it."""
    MM = os.path.join(HOST, "find_mems_memory_mapped"); CH = os.path.join(HOST, "find_mems_chunked")
    if case == "xy_bytecode":
        ri, tg, layout, text = datasets.gold("xy.ri"), datasets.gold("xy_bidirectional_compressed.tags"), 1, datasets.gold("contigs_xy")
    else:
        s = datasets.synth_index(); ri, tg, layout, text = s["ri"], s["tags0"], 0, s["text"]
    reads, off = datasets.sample_reads(text, 400, 150, 0.02, seed=5)
    rs = [reads[int(off[i]):int(off[i + 1])].tobytes() for i in range(400)]
    (tmp_path / "x.ri").write_bytes(ri); (tmp_path / "x.tags").write_bytes(tg); (tmp_path / "reads.txt").write_bytes(b"\n".join(rs) + b"\n")
    want = orc.Index(ri, tg, layout).find_mems(reads, off, 20, 1)["text"]
    base = [str(tmp_path / "x.ri"), str(tmp_path / "x.tags"), str(tmp_path / "reads.txt"), "20", "1"]
    img = str(tmp_path / "x.ri.pgimg")
    p = subprocess.run([MM] + base, capture_output=True)
    assert p.returncode == 0, p.stderr.decode()
    assert strip_timing(p.stdout) == want and b"Flattening the index into" in p.stderr and os.path.exists(img)
    tail = p.stdout[p.stdout.rfind(b"\nTotal time for finding all MEMs:"):].split(b"\n")
    assert tail[3].startswith(b"Memory-mapped indices used: ") and tail[3].endswith(b" GB") and tail[4:] == [b""]
    stamp = os.path.getmtime(img)
    p = subprocess.run([MM] + base + ["--batch", "64", "--verify-image"], capture_output=True)
    assert p.returncode == 0 and strip_timing(p.stdout) == want and b"Using the cached image" in p.stderr and os.path.getmtime(img) == stamp
    for extra, mb in ([], 2048), (["1"], 1), (["3", "--batch", "77"], 3):
        p = subprocess.run([CH] + base + extra, capture_output=True)
        assert p.returncode == 0, p.stderr.decode()
        assert strip_timing(p.stdout) == want and b"Using the cached image" in p.stderr
        tail = p.stdout[p.stdout.rfind(b"\nTotal time for finding all MEMs:"):].split(b"\n")
        assert tail[3].startswith(b"Chunked memory-mapped indices used: ") and tail[4] == b"Chunk size: %d MB" % mb and tail[5:] == [b""]
        assert b"Chunk size: %d MB" % mb in p.stderr
    # a changed source file makes the image stale
    s2 = datasets.synth_index(seed=8) if case == "synth_compact" else None
    if s2:
        (tmp_path / "x.ri").write_bytes(s2["ri"]); (tmp_path / "x.tags").write_bytes(s2["tags0"])
        r2, o2 = datasets.sample_reads(s2["text"], 100, 150, 0.02, seed=6)
        (tmp_path / "reads.txt").write_bytes(b"\n".join(r2[int(o2[i]):int(o2[i + 1])].tobytes() for i in range(100)) + b"\n")
        p = subprocess.run([CH] + base, capture_output=True)
        assert p.returncode == 0 and b"Flattening the index into" in p.stderr
        assert strip_timing(p.stdout) == orc.Index(s2["ri"], s2["tags0"], 0).find_mems(r2, o2, 20, 1)["text"]
    p = subprocess.run([MM, str(tmp_path / "nope.ri")] + base[1:], capture_output=True)
    assert p.returncode != 0 and b"Failed to open memory-mapped indices" in p.stderr
