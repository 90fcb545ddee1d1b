"""Pins oracle/ (the CPU restatement every GPU parity test is checked against) to the REFERENCE'S OWN CODE.

oracle/_ref/libpgiref.so is built by oracle/Makefile from the reference's unmodified src/r-index.cpp and
src/tag_arrays.cpp plus find_mems_function / find_all_mems (algorithm.hpp:644-757) and the main loops of find_mems.cpp /
query_tags.cpp cut out by line range, against stand-in SDSL / GBWT headers (oracle/ref_shim). Where it exists (the build
container, and the GPU box, which receives it prebuilt) the restatement is compared with it call by call; everywhere,
the restatement is compared with the outputs the reference's code printed, committed as tests/golden/ref_find_mems.json
by tests/golden/make_ref_golden.py."""
import hashlib, json, os
import numpy as np, pytest
import orc, datasets, refcases

GOLD = json.load(open(os.path.join(datasets.GOLD, "ref_find_mems.json")))
IX = None
def index(name):
    global IX
    if IX is None: IX = refcases.indexes()
    return IX[name]
def sha(b): return hashlib.sha256(b).hexdigest()
needs_ref = pytest.mark.skipif(not orc.ref_available(), reason="oracle/_ref/libpgiref.so is not built (no /root/reference on this machine)")

def oracle_text(ix, rs, ml, mo):
    o = orc.Index(ix["ri"], ix["tags"], ix["layout"])
    reads, off = orc.pack_reads(rs)
    return o.find_mems(reads, off, ml, mo)["text"]

def test_golden_covers_enough():
    assert GOLD["_reads_total"] >= 10000 and len(GOLD["find_mems"]) == len(refcases.CASES)
    assert any(v["verbatim_main_loop"] for v in GOLD["find_mems"].values())

@pytest.mark.parametrize("case", refcases.CASES, ids=refcases.case_key)
def test_oracle_stdout_equals_reference_golden(case):
    """Restatement == what the reference's own main loop printed (MEM boundaries, BWT intervals, tag lists, grammar)."""
    ix = index(case[0]); g = GOLD["find_mems"][refcases.case_key(case)]
    rs = refcases.reads_of(ix, case[1]); assert len(rs) == g["n_reads"]
    text = oracle_text(ix, rs, case[2], case[3])
    assert len(text) == g["stdout_bytes"] and sha(text) == g["stdout_sha256"]
    if "stdout" in g: assert text.decode() == g["stdout"]

@pytest.mark.parametrize("name", sorted(GOLD["extend_rank"]))
def test_oracle_extend_rank_equal_reference_golden(name):
    from importlib import import_module
    mk = import_module("golden.make_ref_golden") if False else None
    ix = index(name); o = orc.Index(ix["ri"]); g = GOLD["extend_rank"][name]
    iv, sym, d, pos = _vectors(o.n)
    assert o.n == g["n"]
    assert sha(o.extend_batch(iv, sym, d).tobytes()) == g["extend_sha256"]
    assert sha(o.rank_batch(pos).tobytes()) == g["rank_sha256"]

def _vectors(n, seed=3, N=20000):          # the generator of tests/golden/make_ref_golden.py::extend_vectors
    rng = np.random.default_rng(seed)
    iv = np.zeros((N, 3), dtype=np.uint64)
    k = rng.integers(0, n, N); sz = np.minimum(rng.integers(0, 300, N), n - k)
    iv[:, 0] = k; iv[:, 1] = rng.integers(0, n, N); iv[:, 2] = sz
    iv[:200] = [0, 0, n]
    sym = rng.choice(np.frombuffer(b"ACGTN\n\0acgtnX", dtype=np.uint8), N); d = rng.integers(0, 2, N).astype(np.uint8)
    pos = np.concatenate([rng.integers(0, n + 1, 5000), [0, n, n - 1, 1]]).astype(np.uint64)
    return iv, sym, d, pos

@pytest.mark.parametrize("name", sorted(GOLD["query_tags"]))
def test_oracle_query_tags_stdout_equals_reference_golden(name):
    ix = index(name); o = orc.Index(ix["ri"], ix["tags"], 0); g = GOLD["query_tags"][name]
    rs = refcases.reads_of(ix, ("sample", 400, 40, 0.01, 12)); reads, off = orc.pack_reads(rs)
    text = o.query_tags(reads, off)["text"]
    assert len(text) == g["stdout_bytes"] and sha(text) == g["stdout_sha256"]

# ---------------------------------------------------------------------------------------- live, where the reference code is built
@needs_ref
def test_reference_code_reads_shipped_files_to_eof_and_reproduces_known_answers():
    ri = datasets.gold("xy.ri"); tg = datasets.gold("xy_bidirectional_compressed.tags")
    R = orc.Index(ri, tg, 1, impl="ref")
    assert R.ri_consumed == len(ri) and R.tags_consumed == len(tg) and R.n == 8022 and R.C == [0, 8, 2317, 4015, 5713]
    # SURVEY.md App. G known answers (self-derived at survey time; here produced by the reference's own backward_extend)
    for pat, want in ((b"T", (5713, 8, 8)), (b"A", (8, 3420, 2309)), (b"C", (2317, 1714, 1698)), (b"G", (4015, 16, 1698)),
                      (b"ACGA", (969, 3589, 4)), (b"GGCA", (5071, 4021, 21)), (b"ACGT", (0, 0, 0))):
        assert R.search(pat) == want
    # SURVEY.md App. C worked example
    t = R.ref_text(datasets.gold("reads.txt"), 5, 1, verbatim=False).decode()
    assert "Seq: 2\nMEM START: 3, MEM END: 10 BWT START: 1381 SIZE: 4\nNumber of unique positions: 3\n62467, 128008, 203779, \n" in t

@needs_ref
def test_reference_constructor_and_serialisers_regenerate_shipped_ri(tmp_path):
    """FastLocate(rlbwt) + serialize through the stand-in SDSL writes the shipped legacy .ri files byte for byte (pins the
    shim's layouts), and serialize_encoded writes the same bytes as our prep tool (pins prep/ and the encoded layout)."""
    prep = refcases.load_package().prep
    for rl, ri in (("contigs_xy.rl_bwt", "xy.ri"), ("contigs_XY.rl_bwt", "two_contig_xy.ri")):
        p = tmp_path / rl; p.write_bytes(datasets.gold(rl))
        assert orc.ref_build_rindex(str(p), encoded=False) == datasets.gold(ri)
        assert orc.ref_build_rindex(str(p), encoded=True) == prep.rindex_from_rlbwt(datasets.gold(rl), encoded=True, with_samples=True)[0]
    s = datasets.synth_index(encoded=True)
    p = tmp_path / "synth.rl_bwt"; p.write_bytes(s["rl_bwt"])
    assert orc.ref_build_rindex(str(p), encoded=True) == prep.rindex_from_rlbwt(s["rl_bwt"], encoded=True, with_samples=True)[0]

@needs_ref
@pytest.mark.parametrize("name", ["xy_legacy_bytecode", "two_contig_legacy", "x_encoded_sigma5", "synth_enc_compact", "synth_leg_compact"])
def test_oracle_extend_rank_equal_reference_live(name):
    ix = index(name); o = orc.Index(ix["ri"]); R = orc.Index(ix["ri"], impl="ref")
    iv, sym, d, pos = _vectors(o.n, seed=17, N=50000)
    assert np.array_equal(o.extend_batch(iv, sym, d), R.extend_batch(iv, sym, d))
    allpos = np.arange(0, min(o.n, 20000) + 1, dtype=np.uint64)
    assert np.array_equal(o.rank_batch(allpos), R.rank_batch(allpos))

@needs_ref
@pytest.mark.parametrize("case", refcases.CASES[2:], ids=refcases.case_key)
def test_oracle_find_mems_arrays_equal_reference_live(case):
    """MEM records, tag lists and number_of_runs as arrays (what the GPU tests compare), reference code vs restatement."""
    ix = index(case[0]); rs = refcases.reads_of(ix, case[1]); reads, off = orc.pack_reads(rs)
    o = orc.Index(ix["ri"], ix["tags"], ix["layout"]); R = orc.Index(ix["ri"], ix["tags"], ix["layout"], impl="ref")
    a = o.find_mems(reads, off, case[2], case[3], text=False); b = R.find_mems(reads, off, case[2], case[3], text=False, threads=2)
    for k in ("n_mems", "n_tags"): assert a[k] == b[k]
    for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(a[k], b[k]), k

@needs_ref
@pytest.mark.parametrize("name", ["xy_legacy_bytecode", "xy_legacy_compact", "synth_enc_compact", "synth_enc_bytecode"])
def test_oracle_tag_queries_equal_reference_live(name):
    """Every start position with short intervals (hits first % 10 == 0) + random long ones; intervals stay
    inside the array (beyond it the reference reads out of bounds)."""
    ix = index(name); o = orc.Index(ix["ri"], ix["tags"], ix["layout"]); R = orc.Index(ix["ri"], ix["tags"], ix["layout"], impl="ref")
    n = o.n; rng = np.random.default_rng(5)
    hi = n - 60                                            # keep clear of the last decade of runs (undefined there)
    qs = [(s, s + 3) for s in range(0, min(hi, 3000))] + [(int(s), int(min(hi, s + l))) for s, l in zip(rng.integers(0, hi, 1500), rng.integers(1, 2500, 1500))]
    bad = 0
    for s, e in qs:
        a = o.tag_query(s, e); b = R.tag_query(s, e)
        assert a[0] == b[0] and np.array_equal(a[1], b[1]), (s, e)

@needs_ref
def test_reference_threads_do_not_change_results():
    ix = index("synth_enc_compact"); R = orc.Index(ix["ri"], ix["tags"], 0, impl="ref")
    reads, off = datasets.sample_reads(ix["text"], 500, 150, 0.01, seed=1)
    a = R.find_mems(reads, off, 20, 1, text=False, threads=1); b = R.find_mems(reads, off, 20, 1, text=False, threads=4)
    assert np.array_equal(a["mems"], b["mems"]) and np.array_equal(a["tags"], b["tags"])
    dt, nm = R.ref_time(reads, off, 20, 1, threads=2); assert nm == a["n_mems"] and dt > 0
