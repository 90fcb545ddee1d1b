"""CPU: the oracle (oracle/, restatement of the reference's FastLocate / TagArray / find_all_mems) against the
reference's shipped fixtures (tests/golden/, copied from /root/reference/test_data by make_golden.py) and the
known-answer vectors of SURVEY.md App. C / G. No GPU needed."""
import hashlib, json, os
import numpy as np, pytest
import orc, datasets
from conftest import load_package

def rl_bwt_expand(buf: bytes) -> bytes:
    """deps/grlBWT/include/bwt_io.h:40-58: u64 sb, u64 fb, then (symbol: sb bytes, length: fb bytes) records."""
    sb, fb = (int(x) for x in np.frombuffer(buf[:16], dtype=np.uint64))
    rec = np.frombuffer(buf[16:], dtype=np.uint8).reshape(-1, sb + fb)
    lens = np.zeros(len(rec), dtype=np.int64)
    for i in range(fb):
        lens |= rec[:, sb + i].astype(np.int64) << (8 * i)
    return np.repeat(rec[:, 0], lens).tobytes()

def naive_bwt(text: bytes) -> bytes:
    """tests/test_rindex.cpp:35-99 of the reference: multi-string BWT with sentinels $1 < $2 < ... < 'A', rotations."""
    arr = []; k = 0
    for ch in text:
        if ch == 10: arr.append(k); k += 1
        else: arr.append(ch + 1000)
    n = len(arr); dbl = arr + arr
    order = sorted(range(n), key=lambda i: dbl[i:i + n])
    return bytes(text[(i - 1) % n] for i in order)

def test_golden_manifest():
    man = json.load(open(os.path.join(datasets.GOLD, "MANIFEST.json")))
    assert len(man) >= 17
    for name, m in man.items():
        b = datasets.gold(name)
        assert len(b) == m["bytes"] and hashlib.sha256(b).hexdigest() == m["sha256"], name

@pytest.mark.parametrize("ri,n,sigma,C,blocks", [
    ("xy.ri", 8022, 5, [0, 8, 2317, 4015, 5713], 162),
    ("two_contig_xy.ri", 4011, 5, [0, 4, 1016, 1912, 2714], 81)])
def test_shipped_ri_parsed_to_eof(ri, n, sigma, C, blocks):
    """the reader consumes the shipped legacy FastLocate files to the last byte."""
    b = datasets.gold(ri); o = orc.Index(b)
    assert o.ri_consumed == len(b)
    assert (o.n, o.sigma, o.C) == (n, sigma, C)
    assert o.info[2] == 0 and o.info[4] == blocks                # legacy layout; ones in blocks_start_pos
    assert o.sym_map[ord("T")] == 4 and o.sym_map[ord("N")] == 0 and o.sym_map[0] == 0   # sigma = 5: no N

def test_shipped_tags_parsed_to_eof():
    b = datasets.gold("xy_bidirectional_compressed.tags"); o = orc.Index(datasets.gold("xy.ri"), b, 1)
    assert o.tags_consumed == len(b)
    assert o.info[10] == 24087 and o.info[11] == 604 and o.info[12] == 6031 and o.info[13] == 8038

def test_bad_files_rejected():
    with pytest.raises(RuntimeError): orc.Index(b"\0" * 64)
    with pytest.raises(RuntimeError): orc.Index(datasets.gold("xy.ri")[:1000])

def test_known_answer_intervals():
    """SURVEY.md App. G.2 (reference semantics incl. the sigma=5 'T' quirk) on bidirectional_test/xy.ri."""
    o = orc.Index(datasets.gold("xy.ri"))
    want = {b"T": (5713, 8, 8), b"A": (8, 3420, 2309), b"C": (2317, 1714, 1698), b"G": (4015, 16, 1698),
            b"ACGA": (969, 3589, 4), b"GGCA": (5071, 4021, 21), b"ACGT": (0, 0, 0)}
    for p, iv in want.items():
        assert o.search(p) == iv, p

def test_worked_example_stdout():
    """reads.txt through find_mems 5 1 with the shipped ByteCode tags."""
    o = orc.Index(datasets.gold("xy.ri"), datasets.gold("xy_bidirectional_compressed.tags"), 1)
    reads = datasets.gold("reads.txt").split()
    assert reads[1] == b"GGTAGCCATGCT"
    buf, off = orc.pack_reads(reads)
    R = o.find_mems(buf, off, 5, 1)
    blocks = R["text"].decode().split("\n\n")
    assert blocks[0] == "Seq: 1"
    assert blocks[1] == ("Seq: 2\nMEM START: 3, MEM END: 10 BWT START: 1381 SIZE: 4\n"
                         "Number of unique positions: 3\n62467, 128008, 203779, ")
    assert R["mem_off"].tolist()[:3] == [0, 0, 1] and R["mems"][0].tolist() == [3, 10, 1381, 4]
    assert R["n_runs"][0] >= 3

def test_shipped_rl_bwt_match_naive_bwt():
    """The BWT convention (reference tests/test_rindex.cpp:35-99) reproduces the shipped grlBWT files run for run."""
    for txt, rl in [("contigs_xy", "contigs_xy.rl_bwt"), ("contigs_x", "contigs_x.rl_bwt"), ("contigs_XY.txt", "contigs_XY.rl_bwt"),
                    ("x.newline_separated", "x.rl_bwt"), ("med_test.txt", "med_test.rl_bwt"), ("small_test.txt", "small_test.rl_bwt")]:
        t = datasets.gold(txt)
        if not t.endswith(b"\n"): t += b"\n"
        assert rl_bwt_expand(datasets.gold(rl)) == naive_bwt(t), txt

def test_prep_regenerates_shipped_files():
    """prep/ (our restated FastLocate(string) constructor + serialize) rebuilds both shipped .ri files byte for byte,
    and its BWT builder agrees with the shipped .rl_bwt files."""
    prep = load_package().prep
    for rl, ri in [("contigs_xy.rl_bwt", "xy.ri"), ("contigs_XY.rl_bwt", "two_contig_xy.ri")]:
        mine, info = prep.rindex_from_rlbwt(datasets.gold(rl), encoded=False, with_samples=True)
        assert mine == datasets.gold(ri), ri
    for txt, rl in [("contigs_xy", "contigs_xy.rl_bwt"), ("x.newline_separated", "x.rl_bwt"), ("med_test.txt", "med_test.rl_bwt")]:
        t = datasets.gold(txt)
        if not t.endswith(b"\n"): t += b"\n"
        assert rl_bwt_expand(prep.bwt_from_text(t)) == rl_bwt_expand(datasets.gold(rl)), txt

def test_encoded_equals_legacy_on_shipped_bwt():
    """serialize_encoded round trip (reference tests Locate_*_encoded): the encoded file of the same BWT answers
    every sigma=6-safe query like the legacy one. On sigma=5 the 'T' slot differs BY DESIGN of the reference
    (legacy: block-start endmarker count, encoded: endmarker rank) — checked explicitly."""
    prep = load_package().prep
    enc, _ = prep.rindex_from_rlbwt(datasets.gold("contigs_xy.rl_bwt"), encoded=True, with_samples=True)
    a = orc.Index(datasets.gold("xy.ri")); b = orc.Index(enc)
    assert b.info[2] == 1 and b.ri_consumed == len(enc)
    pos = np.arange(0, a.n + 1, dtype=np.uint64)
    ra, rb = a.rank_batch(pos), b.rank_batch(pos)
    assert np.array_equal(ra[:, :4], rb[:, :4])                 # '\n', A, C, G: true ranks in both
    assert np.array_equal(rb[:, 4], rb[:, 0])                   # encoded: slot 4 = endmarker rank
    assert (ra[:, 4] <= ra[:, 0]).all() and not np.array_equal(ra[:, 4], ra[:, 0])   # legacy: block-start value

def test_x_index_rebuilt_from_rl_bwt():
    """BASELINE config 1: test_data/x.giraffe.ri is not a FastLocate file; the index is rebuilt
    from x.rl_bwt. Forward-only text (3 sequences), sigma = 5."""
    prep = load_package().prep
    ri, info = prep.rindex_from_rlbwt(datasets.gold("x.rl_bwt"), encoded=True, with_samples=True)
    assert info["n"] == 3012 and info["n_seq"] == 3 and info["sigma"] == 5
    o = orc.Index(ri)
    text = datasets.gold("x.newline_separated")
    for pat in (b"GA", b"CAG", b"AAGC"):                        # T-free patterns count true occurrences
        cnt = sum(1 for i in range(len(text)) if text.startswith(pat, i))
        assert o.search(pat)[2] == cnt, pat

def naive_suffix_order(text: bytes):
    """(sequence id, offset in the sequence) of every BWT row: the rotation sort of tests/test_rindex.cpp:35-99."""
    arr, seq, offs = [], [], []; k = sid = o = 0
    for ch in text:
        seq.append(sid); offs.append(o)
        if ch == 10: arr.append(k); k += 1; sid += 1; o = 0
        else: arr.append(ch + 1000); o += 1
    n = len(arr); dbl = arr + arr
    order = sorted(range(n), key=lambda i: dbl[i:i + n])
    return np.array([seq[i] for i in order], dtype=np.uint64), np.array([offs[i] for i in order], dtype=np.uint64)

@pytest.mark.parametrize("ri,text", [("xy.ri", "contigs_xy"), ("two_contig_xy.ri", "contigs_XY.txt")])
def test_decompress_da_of_shipped_ri_equals_naive_document_array(ri, text):
    """The reference's own locate tests (tests/test_rindex.cpp:103-244) assert decompressDA() == the document array of a
    naive rotation sort. Same assertion for the oracle's getSample / locateNext / seqId on the SHIPPED .ri files and texts
    (plus the offsets, which the reference's test does not look at): pins the locate restatement."""
    o = orc.Index(datasets.gold(ri)); ml = o.info[5]
    sa = o.decompress_sa(); ids, offs = naive_suffix_order(datasets.gold(text))
    assert np.array_equal(sa // np.uint64(ml), ids) and np.array_equal(sa % np.uint64(ml), offs)
    off, uids, rows = o.locate_batch(np.arange(o.n), np.arange(o.n), want_sa=True)      # every row through locate_encoded's run lookup
    assert np.array_equal(rows, sa) and np.array_equal(uids, ids)
    off, uids = o.locate_batch([0, 7, 100, 5, 9], [o.n - 1, 7, 180, 4, 8])
    assert np.array_equal(uids[off[0]:off[1]], np.unique(ids)) and uids[off[1]] == ids[7] and off[4] == off[3]
    assert np.array_equal(uids[off[2]:off[3]], np.unique(ids[100:181]))

def test_build_rindex_cli_regenerates_shipped_files(tmp_path):
    """prep/build_rindex (src/build_rindex.cpp:11-20: FastLocate(rlbwt) + serialize_encoded to stdout; --legacy =
    FastLocate::serialize): the legacy output IS the shipped .ri, byte for byte; the encoded output parses to EOF and
    holds the same BWT."""
    import subprocess
    from conftest import ROOT
    exe = os.path.join(ROOT, "pangenome-index_b200", "prep", "build_rindex")
    for rl, ri in [("contigs_xy.rl_bwt", "xy.ri"), ("contigs_XY.rl_bwt", "two_contig_xy.ri")]:
        src = os.path.join(datasets.GOLD, rl)
        p = subprocess.run([exe, src, "--legacy"], capture_output=True)
        assert p.returncode == 0 and p.stdout == datasets.gold(ri), ri
        p = subprocess.run([exe, src], capture_output=True)
        assert p.returncode == 0 and b"ENCODED_BLOCKS" in p.stderr
        a = orc.Index(datasets.gold(ri)); b = orc.Index(p.stdout)
        assert b.info[2] == 1 and b.ri_consumed == len(p.stdout) and b.n == a.n
        pos = np.arange(0, a.n + 1, dtype=np.uint64)
        assert np.array_equal(a.rank_batch(pos)[:, :4], b.rank_batch(pos)[:, :4])
    p = subprocess.run([exe, str(tmp_path / "missing.rl_bwt")], capture_output=True)
    assert p.returncode != 0 and b"cannot open" in p.stderr and p.stdout == b""
    p = subprocess.run([exe], capture_output=True)
    assert p.returncode != 0 and b"Usage" in p.stderr
