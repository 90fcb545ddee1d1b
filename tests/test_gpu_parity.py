"""GPU parity: the CUDA path (through the C ABI of include/pgi.h) against the CPU oracle, bit-exact."""
import numpy as np, pytest
import orc, datasets
from conftest import load_package

pytestmark = pytest.mark.gpu

@pytest.fixture(scope="module")
def capi():
    c = load_package().capi
    assert c.lib().pgi_device_count() > 0, "no CUDA device visible: the product has no CPU path"
    return c

def cases():
    xy = dict(ri=datasets.gold("xy.ri"), tags=datasets.gold("xy_bidirectional_compressed.tags"), layout=1, text=datasets.gold("contigs_xy"))
    two = dict(ri=datasets.gold("two_contig_xy.ri"), tags=None, layout=0, text=datasets.gold("contigs_XY.txt"))
    out = {"xy_legacy_bytecode": xy, "two_contig_legacy": two}
    for enc in (True, False):
        s = datasets.synth_index(encoded=enc)
        out["synth_%s_compact" % ("enc" if enc else "leg")] = dict(ri=s["ri"], tags=s["tags0"], layout=0, text=s["text"])
    s = datasets.synth_index(encoded=True)
    out["synth_enc_bytecode"] = dict(ri=s["ri"], tags=s["tags1"], layout=1, text=s["text"])
    import refcases
    out["x_encoded_sigma5"] = refcases.x_index()              # BASELINE config 1: encoded, forward-only, sigma = 5
    return out

CASES = None
def get_case(name):
    global CASES
    if CASES is None: CASES = cases()
    return CASES[name]
NAMES = ["xy_legacy_bytecode", "two_contig_legacy", "synth_enc_compact", "synth_leg_compact", "synth_enc_bytecode", "x_encoded_sigma5"]

def compare(g, o, with_tags):
    assert g["n_mems"] == o["n_mems"]
    assert np.array_equal(g["mem_off"], o["mem_off"])
    assert np.array_equal(g["mems"], o["mems"])
    if with_tags:
        assert np.array_equal(g["n_runs"], o["n_runs"])
        assert np.array_equal(g["tag_off"], o["tag_off"])
        assert np.array_equal(g["tags"], o["tags"])
    assert g["text"] == o["text"]

# ---- against the REFERENCE'S OWN CODE: its stdout on the cases of tests/refcases.py, committed as tests/golden/ref_find_mems.json
# by tests/golden/make_ref_golden.py (oracle/_ref/libpgiref.so = src/r-index.cpp + src/tag_arrays.cpp + find_all_mems + main loop)
import hashlib, json, os, refcases
REF_GOLD = json.load(open(os.path.join(datasets.GOLD, "ref_find_mems.json")))
_REF_IX = None
def ref_index(name):
    global _REF_IX
    if _REF_IX is None: _REF_IX = refcases.indexes()
    return _REF_IX[name]

@pytest.mark.parametrize("case", refcases.CASES, ids=refcases.case_key)
def test_gpu_stdout_equals_reference_golden(capi, case):
    ix = ref_index(case[0]); gold = REF_GOLD["find_mems"][refcases.case_key(case)]
    rs = refcases.reads_of(ix, case[1]); assert len(rs) == gold["n_reads"]
    reads, off = orc.pack_reads(rs)
    g = capi.Index(ix["ri"], ix["tags"], ix["layout"])
    text = g.find_mems(reads, off, case[2], case[3])["text"]
    assert len(text) == gold["stdout_bytes"] and hashlib.sha256(text).hexdigest() == gold["stdout_sha256"]
    if orc.ref_available():                                   # the prebuilt reference library travels to the GPU box: compare live as well
        R = orc.Index(ix["ri"], ix["tags"], ix["layout"], impl="ref")
        assert text == R.ref_text(b"\n".join(rs) + b"\n", case[2], case[3])

@pytest.mark.parametrize("name", sorted(REF_GOLD["extend_rank"]))
def test_gpu_extend_rank_equal_reference_golden(capi, name):
    from test_ref_pin import _vectors
    ix = ref_index(name); g = capi.Index(ix["ri"]); gold = REF_GOLD["extend_rank"][name]
    iv, sym, d, pos = _vectors(g.info["n"])
    assert hashlib.sha256(g.extend_batch(iv, sym, d).tobytes()).hexdigest() == gold["extend_sha256"]
    assert hashlib.sha256(g.rank_batch(pos).tobytes()).hexdigest() == gold["rank_sha256"]

@pytest.mark.parametrize("name", sorted(REF_GOLD["query_tags"]))
def test_gpu_query_tags_stdout_equals_reference_golden(capi, name):
    ix = ref_index(name); g = capi.Index(ix["ri"], ix["tags"], 0); gold = REF_GOLD["query_tags"][name]
    reads, off = orc.pack_reads(refcases.reads_of(ix, ("sample", 400, 40, 0.01, 12)))
    text = g.query_tags(reads, off)["text"]
    assert len(text) == gold["stdout_bytes"] and hashlib.sha256(text).hexdigest() == gold["stdout_sha256"]

@pytest.mark.parametrize("name", NAMES)
def test_extend_and_rank(capi, name):
    c = get_case(name)
    o = orc.Index(c["ri"]); g = capi.Index(c["ri"])
    rng = np.random.default_rng(3); N = 50000; n = o.n
    iv = np.zeros((N, 3), dtype=np.uint64)
    k = rng.integers(0, n, N); sz = np.minimum(rng.integers(0, 300, N), n - k)
    iv[:, 0] = k; iv[:, 1] = rng.integers(0, n, N); iv[:, 2] = sz
    iv[:200] = [0, 0, n]
    iv[200:300, 0] = rng.integers(0, 2 * n, 100)             # positions beyond n (defined UB: read as n)
    sym = rng.choice(np.frombuffer(b"ACGTN\n\0acgtnX\xff", dtype=np.uint8), N); d = rng.integers(0, 2, N).astype(np.uint8)
    assert np.array_equal(g.extend_batch(iv, sym, d), o.extend_batch(iv, sym, d))
    pos = np.concatenate([rng.integers(0, n + 1, 20000), [0, n, n - 1, 1]]).astype(np.uint64)
    assert np.array_equal(g.rank_batch(pos), o.rank_batch(pos))

@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("params", [(20, 1), (5, 1), (1, 1), (10, 3), (31, 2)])
def test_find_mems(capi, name, params):
    c = get_case(name); ml, mo = params
    o = orc.Index(c["ri"], c["tags"], c["layout"]); g = capi.Index(c["ri"], c["tags"], c["layout"])
    nreads = 3000 if ml >= 10 else 400
    reads, off = datasets.sample_reads(c["text"], nreads, 150, 0.01, seed=ml * 7 + mo)
    R = o.find_mems(reads, off, ml, mo, count_work=True); G = g.find_mems(reads, off, ml, mo)
    compare(G, R, c["tags"] is not None)
    assert G["extensions"] == R["counters"]["E"]

@pytest.mark.parametrize("name", ["xy_legacy_bytecode", "synth_enc_compact"])
def test_ragged_and_edge_reads(capi, name):
    c = get_case(name)
    o = orc.Index(c["ri"], c["tags"], c["layout"]); g = capi.Index(c["ri"], c["tags"], c["layout"])
    rs = datasets.ragged_reads(c["text"], seed=5)
    reads, off = orc.pack_reads(rs)
    for ml, mo in [(20, 1), (3, 1), (1, 2)]:
        compare(g.find_mems(reads, off, ml, mo), o.find_mems(reads, off, ml, mo), True)
    e = np.zeros(0, dtype=np.uint8)                           # empty batch
    G = g.find_mems(e, np.zeros(1, dtype=np.uint64), 20, 1)
    assert G["n_mems"] == 0 and G["n_reads"] == 0

@pytest.mark.parametrize("name", ["xy_legacy_bytecode", "synth_enc_compact", "synth_enc_bytecode"])
def test_tag_query_batch(capi, name):
    c = get_case(name)
    o = orc.Index(c["ri"], c["tags"], c["layout"]); g = capi.Index(c["ri"], c["tags"], c["layout"])
    rng = np.random.default_rng(11); N = 4000; n = o.n
    st = rng.integers(0, n, N).astype(np.uint64)
    ln = np.where(rng.random(N) < 0.9, rng.integers(1, 40, N), rng.integers(40, 3000, N)).astype(np.uint64)
    en = np.minimum(st + ln - 1, np.uint64(n + 40))          # a few intervals run past the end of the array
    G = g.tag_query_batch(st, en)
    for i in range(N):
        nr, tv = o.tag_query(int(st[i]), int(en[i]))
        assert G["n_runs"][i] == nr, i
        a, b = G["tag_off"][i], G["tag_off"][i + 1]
        assert np.array_equal(G["tags"][a:b], tv), i

def test_first_mod_10_quirk(capi):
    """intervals whose first run index is a multiple of 10 start one run late."""
    c = get_case("synth_enc_compact"); o = orc.Index(c["ri"], c["tags"], 0); g = capi.Index(c["ri"], c["tags"], 0)
    st = np.arange(0, min(o.n, 4000), dtype=np.uint64); en = st + np.uint64(3)
    G = g.tag_query_batch(st, en)
    for i in range(len(st)):
        nr, tv = o.tag_query(int(st[i]), int(en[i]))
        a, b = G["tag_off"][i], G["tag_off"][i + 1]
        assert G["n_runs"][i] == nr and np.array_equal(G["tags"][a:b], tv)

def test_clone_and_batch_api(capi):
    c = get_case("synth_enc_compact"); o = orc.Index(c["ri"], c["tags"], 0); g = capi.Index(c["ri"], c["tags"], 0)
    g2 = g.clone_to(0)
    reads, off = datasets.sample_reads(c["text"], 2000, 150, 0.02, seed=9)
    R = o.find_mems(reads, off, 20, 1)
    compare(g2.find_mems(reads, off, 20, 1), R, True)
    b = capi.Batch(g2)
    for _ in range(2):                                        # context reuse
        b.upload(reads.ctypes.data, off.ctypes.data, len(off) - 1); b.run(20, 1)
        G = b.fetch(text=True); compare(G, R, True)
    ms = b.stage_ms(); assert ms[6] >= 4 and ms[5] > 0

def test_large_batch_properties(capi):
    """Many reads; size-independent checks (every MEM is >= min_len, intervals non-empty, offsets monotone,
    tags sorted unique) plus oracle equality on a slice."""
    s = datasets.synth_index(ref_len=20000, n_hap=8, seed=21, encoded=True)
    g = capi.Index(s["ri"], s["tags0"], 0); o = orc.Index(s["ri"], s["tags0"], 0)
    reads, off = datasets.sample_reads(s["text"], 200000, 150, 0.01, seed=4)
    G = g.find_mems(reads, off, 20, 1, text=False)
    m = G["mems"]
    assert (m[:, 1] - m[:, 0] >= 20).all() and (m[:, 3].astype(np.int64) > 0).all() and (m[:, 1] <= 150).all()
    assert (np.diff(G["mem_off"].astype(np.int64)) >= 0).all() and (np.diff(G["tag_off"].astype(np.int64)) >= 0).all()
    for i in np.random.default_rng(0).integers(0, G["n_mems"], 2000):
        t = G["tags"][G["tag_off"][i]:G["tag_off"][i + 1]]
        assert (np.diff(t.astype(np.int64)) > 0).all()
    k = 5000
    R = o.find_mems(reads[: k * 150], off[: k + 1], 20, 1)
    assert np.array_equal(G["mems"][: R["n_mems"]], R["mems"]) and np.array_equal(G["tags"][: R["n_tags"]], R["tags"])

def test_large_ragged_batch_and_concurrent_callers(capi):
    """A large ragged batch through pgi_find_mems equals the explicit batch API bit for bit (and the oracle on a slice);
    several host threads calling pgi_find_mems on one index at the same time (the batched driver's pattern: the copies of
    one call overlap the kernels of another) all get the same answer."""
    s = datasets.synth_index(ref_len=20000, n_hap=8, seed=21, encoded=True)
    g = capi.Index(s["ri"], s["tags0"], 0); o = orc.Index(s["ri"], s["tags0"], 0)
    rng = np.random.default_rng(8)
    base, _ = datasets.sample_reads(s["text"], 300000, 150, 0.02, seed=12)
    lens = rng.integers(0, 151, 300000).astype(np.uint64); lens[:5] = 0; lens[-3:] = 0
    off = np.zeros(300001, dtype=np.uint64); off[1:] = np.cumsum(lens)
    reads = np.concatenate([base[i * 150:i * 150 + int(l)] for i, l in enumerate(lens[:2000])] + [np.zeros(0, np.uint8)])
    # full ragged buffer built vectorised: take the first lens[i] bytes of read i
    idx = np.repeat(np.arange(300000, dtype=np.int64) * 150, lens.astype(np.int64)) + (np.arange(int(off[-1]), dtype=np.int64) - np.repeat(off[:-1].astype(np.int64), lens.astype(np.int64)))
    reads = base[idx]
    G = g.find_mems(reads, off, 20, 1, text=False)
    b = capi.Batch(g); b.upload(reads.ctypes.data, off.ctypes.data, 300000); b.run(20, 1); S = b.fetch()
    for k in ("n_mems", "n_tags", "extensions"): assert G[k] == S[k], k
    for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(G[k], S[k]), k
    import threading
    outs = [None] * 3
    def call(i): outs[i] = g.find_mems(reads, off, 20, 1, text=False)
    ths = [threading.Thread(target=call, args=(i,)) for i in range(3)]
    [t.start() for t in ths]; [t.join() for t in ths]
    for o_ in outs:
        for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(G[k], o_[k]), k
    kk = 4000
    R = o.find_mems(reads[: int(off[kk])], off[: kk + 1], 20, 1, text=False)
    assert np.array_equal(G["mems"][: R["n_mems"]], R["mems"]) and np.array_equal(G["tags"][: R["n_tags"]], R["tags"])
    assert np.array_equal(G["mem_off"][: kk + 1], R["mem_off"])

def test_sharded_equals_unsharded(capi):
    """The multi-GPU host logic (sharding.py) on the CUDA path: shards computed separately (as N ranks would, each on its
    replica) and merged equal the unsharded batch; a clone of the index gives the same answers."""
    sharding = load_package().sharding
    c = get_case("synth_enc_compact"); g = capi.Index(c["ri"], c["tags"], 0); g2 = g.clone_to(0)
    rs = datasets.ragged_reads(c["text"], seed=13, n=600); reads, off = orc.pack_reads(rs)
    whole = g.find_mems(reads, off, 12, 1, text=False)
    for world in (2, 8):
        parts = []
        for r in range(world):
            rr, oo, lo, hi = sharding.shard_reads(reads, off, world, r)
            parts.append((g if r % 2 == 0 else g2).find_mems(rr, oo, 12, 1, text=False))
        m = sharding.merge_results(parts)
        for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(m[k], whole[k]), (world, k)
        assert m["extensions"] == whole["extensions"]
    assert sharding.find_mems_sharded(lambda *a: g.find_mems(*a, text=False), reads, off, 12, 1)["n_mems"] == whole["n_mems"]

def query_reads(text, seed, n=3000):
    """Reads for the whole-read search: substrings of every length (most match), a few with one changed base, edge cases."""
    rng = np.random.default_rng(seed)
    lines = [l for l in text.split(b"\n") if len(l) > 100]
    rs = [b"", b"A", b"T", b"N", b"\n", b"a", b"ACGT", bytes([0, 200]), b"GATTACA"]
    for _ in range(n):
        L = lines[int(rng.integers(len(lines)))]; ln = int(rng.integers(1, 120)); st = int(rng.integers(0, len(L) - ln))
        r = bytearray(L[st:st + ln])
        if rng.random() < 0.3: r[int(rng.integers(len(r)))] = b"ACGTN\nx"[int(rng.integers(7))]
        rs.append(bytes(r))
    return rs

def compare_query_tags(G, R, with_tags):
    assert G["n_reads"] == R["n_reads"] and G["n_mems"] == R["n_mems"]
    assert np.array_equal(G["mem_off"], R["mem_off"]) and np.array_equal(G["mems"], R["mems"])
    if with_tags:
        assert np.array_equal(G["n_runs"], R["n_runs"]) and np.array_equal(G["tag_off"], R["tag_off"]) and np.array_equal(G["tags"], R["tags"])
        assert G["text"] == R["text"]

@pytest.mark.parametrize("name", [n for n in NAMES if n != "x_encoded_sigma5"])   # encoded without N: refused (test_query_tags_large_and_refusal)
def test_query_tags(capi, name):
    """count_encoded / count + tag query of the final range (src/query_tags.cpp:87-109), bit-exact."""
    c = get_case(name)
    o = orc.Index(c["ri"], c["tags"], c["layout"]); g = capi.Index(c["ri"], c["tags"], c["layout"])
    reads, off = orc.pack_reads(query_reads(c["text"], seed=17))
    R = o.query_tags(reads, off); G = g.query_tags(reads, off)
    compare_query_tags(G, R, c["tags"] is not None)
    hit = R["mems"][:, 3] > 0
    assert hit.sum() > 1000 and (~hit).sum() > 100
    assert (R["mems"][~hit, 2] == 1).all()                         # the reference's empty range {1, 0}
    assert G["extensions"] > 0
    e = g.query_tags(np.zeros(0, dtype=np.uint8), np.zeros(1, dtype=np.uint64))     # empty batch
    assert e["n_reads"] == 0 and e["n_mems"] == 0
    b = capi.Batch(g)                                            # batch-context form
    b.upload(reads.ctypes.data, off.ctypes.data, len(off) - 1); b.run_query_tags(); S = b.fetch()
    for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(S[k], G[k]), k

def test_query_tags_large_and_refusal(capi):
    """Many reads against plain-text occurrence counts (size-independent property: interval size = number of
    occurrences of the read in the text), the oracle on a slice, and the refusal on an encoded index without N."""
    s = datasets.synth_index(ref_len=20000, n_hap=8, seed=21, encoded=True)
    g = capi.Index(s["ri"], s["tags0"], 0); o = orc.Index(s["ri"], s["tags0"], 0)
    reads, off = datasets.sample_reads(s["text"], 100000, 40, 0.01, seed=6)
    G = g.query_tags(reads, off, text=False)
    k = 4000
    R = o.query_tags(reads[: k * 40], off[: k + 1], text=False)
    assert np.array_equal(G["mems"][:k], R["mems"]) and np.array_equal(G["tags"][: R["n_tags"]], R["tags"])
    text = s["text"]
    for i in np.random.default_rng(1).integers(0, 100000, 300):
        rd = reads[i * 40:(i + 1) * 40].tobytes(); cnt = 0; p = text.find(rd)
        while p >= 0: cnt += 1; p = text.find(rd, p + 1)
        assert int(G["mems"][i, 3]) == cnt, i
    s5 = datasets.synth_index(ref_len=2000, n_hap=2, seed=3, encoded=True, n_N=0)       # sigma 5, encoded
    g5 = capi.Index(s5["ri"])
    with pytest.raises(capi.PgiError):
        g5.query_tags(reads[:400], off[:11])
    r5, o5 = datasets.sample_reads(s5["text"], 500, 150, 0.01, seed=2)                  # find_mems itself works on it
    compare(g5.find_mems(r5, o5, 20, 1), orc.Index(s5["ri"]).find_mems(r5, o5, 20, 1), False)

@pytest.mark.parametrize("shift", ["12", "10", "8", "4"])
def test_crowded_buckets(capi, shift, monkeypatch):
    """Bucket size forced (PGI_BSHIFT) so that most buckets are crowded (8: directory block + cooperatively fetched chunk
    blocks on every trip) or tiny (4): the answers do not depend on the device layout."""
    monkeypatch.setenv("PGI_BSHIFT", shift)
    for name in ("xy_legacy_bytecode", "synth_enc_compact"):
        c = get_case(name)
        g = capi.Index(c["ri"], c["tags"], c["layout"]); o = orc.Index(c["ri"], c["tags"], c["layout"])
        assert g.info["bucket_shift"] == int(shift)
        if shift == "8": assert g.info["n_crowded_buckets"] > g.info["n"] // 256 // 2
        reads, off = datasets.sample_reads(c["text"], 3000, 150, 0.01, seed=23)
        compare(g.find_mems(reads, off, 20, 1), o.find_mems(reads, off, 20, 1), True)
        rs = datasets.ragged_reads(c["text"], seed=6); r2, o2 = orc.pack_reads(rs)
        compare(g.find_mems(r2, o2, 3, 1), o.find_mems(r2, o2, 3, 1), True)
        q, qo = orc.pack_reads(query_reads(c["text"], seed=3, n=1500))
        compare_query_tags(g.query_tags(q, qo), o.query_tags(q, qo), True)
        rng = np.random.default_rng(2); pos = rng.integers(0, o.n + 1, 20000).astype(np.uint64)
        assert np.array_equal(g.rank_batch(pos), o.rank_batch(pos))

@pytest.mark.parametrize("K", ["0", "1", "4", "7", "8", "11", "12"])
def test_seed_table_is_invisible(capi, K, monkeypatch):
    """Seed table (pgi_core.cuh: the first K backward extensions of a walk from the full interval come from ONE lookup),
    K forced through PGI_SEEDK: identical MEMs, tags, text AND extension count for every K, for min_len below / equal /
    above K, min_occ > 1, ragged reads, other bytes, long reads; a cloned index carries the table; an index with a quirk
    slot gets none."""
    monkeypatch.setenv("PGI_SEEDK", K)
    c = get_case("synth_enc_compact")
    g = capi.Index(c["ri"], c["tags"], c["layout"]); o = orc.Index(c["ri"], c["tags"], c["layout"])
    assert g.info["seed_k"] == int(K)
    k = int(K)
    for ml, mo in [(20, 1), (max(1, k), 1), (max(1, k - 1), 1), (k + 1, 2), (12, 40)]:
        reads, off = datasets.sample_reads(c["text"], 4000, 150, 0.02, seed=40 + k + ml)
        G = g.find_mems(reads, off, ml, mo); R = o.find_mems(reads, off, ml, mo, count_work=True)
        compare(G, R, True)
        assert G["extensions"] == R["counters"]["E"], (ml, mo)
        # the lookup counter: none without a table; with one, every lookup stood for k of the algorithm's extensions
        assert G["seed_lookups"] * k <= G["extensions"]
        if k == 0 or ml < k: assert G["seed_lookups"] == 0 or ml < k
        if 0 < k <= ml and mo == 1: assert G["seed_lookups"] > 0
    rs = datasets.ragged_reads(c["text"], seed=8); r2, o2 = orc.pack_reads(rs)
    compare(g.find_mems(r2, o2, 3, 1), o.find_mems(r2, o2, 3, 1), True)
    compare(g.find_mems(r2, o2, 13, 1), o.find_mems(r2, o2, 13, 1), True)
    q, qo = orc.pack_reads(query_reads(c["text"], seed=5, n=800))
    compare_query_tags(g.query_tags(q, qo), o.query_tags(q, qo), True)               # OP 1 never uses the table
    if K == "8":
        s = datasets.synth_index(ref_len=20000, n_hap=8, seed=21, encoded=True)
        gl = capi.Index(s["ri"], s["tags0"], 0); ol = orc.Index(s["ri"], s["tags0"], 0)
        reads, off = datasets.sample_reads(s["text"], 200, 10000, 0.04, seed=33, indel=0.01)
        G = gl.find_mems(reads, off, 20, 1); R = ol.find_mems(reads, off, 20, 1, count_work=True)
        compare(G, R, True); assert G["extensions"] == R["counters"]["E"]
        cl = g.clone_to(0)
        assert cl.info["seed_k"] == 8
        reads, off = datasets.sample_reads(c["text"], 2000, 150, 0.02, seed=77)
        compare(cl.find_mems(reads, off, 20, 1), o.find_mems(reads, off, 20, 1), True)
        x = get_case("xy_legacy_bytecode"); gx = capi.Index(x["ri"], x["tags"], x["layout"])
        assert gx.info["seed_k"] == (0 if gx.info["has_quirk_slot"] else 8)
        reads, off = datasets.sample_reads(x["text"], 2000, 150, 0.02, seed=78)
        compare(gx.find_mems(reads, off, 20, 1), orc.Index(x["ri"], x["tags"], x["layout"]).find_mems(reads, off, 20, 1), True)

def test_partitioned_mem_slots(capi, monkeypatch):
    """MEM slots handed out from 256 partition counters instead of one global counter: forced on for
    every batch size (PGI_SLOT_PARTS=1; by default only batches of 600000 reads and more use them), including batches that
    overflow a partition and are re-run, long reads with many MEMs each, ragged and empty reads; same results as the
    oracle and as the single-counter path."""
    c = get_case("synth_enc_compact")
    g = capi.Index(c["ri"], c["tags"], c["layout"]); o = orc.Index(c["ri"], c["tags"], c["layout"])
    reads, off = datasets.sample_reads(c["text"], 20000, 150, 0.02, seed=91)
    rs = datasets.ragged_reads(c["text"], seed=12); r2, o2 = orc.pack_reads(rs)
    s = datasets.synth_index(ref_len=20000, n_hap=8, seed=21, encoded=True)
    gl = capi.Index(s["ri"], s["tags0"], 0); ol = orc.Index(s["ri"], s["tags0"], 0)
    lr, lo = datasets.sample_reads(s["text"], 200, 10000, 0.04, seed=35, indel=0.01)
    monkeypatch.setenv("PGI_SLOT_PARTS", "0")
    single = g.find_mems(reads, off, 20, 1)
    monkeypatch.setenv("PGI_SLOT_PARTS", "1")
    for ml, mo in [(20, 1), (5, 2)]:
        G = g.find_mems(reads, off, ml, mo); R = o.find_mems(reads, off, ml, mo, count_work=True)
        compare(G, R, True); assert G["extensions"] == R["counters"]["E"]
    parts = g.find_mems(reads, off, 20, 1)
    for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(parts[k], single[k]), k
    compare(g.find_mems(r2, o2, 3, 1), o.find_mems(r2, o2, 3, 1), True)
    e_reads, e_off = orc.pack_reads([b"", b"ACGT", b""])
    compare(g.find_mems(e_reads, e_off, 2, 1), o.find_mems(e_reads, e_off, 2, 1), True)
    G = gl.find_mems(lr, lo, 20, 1); R = ol.find_mems(lr, lo, 20, 1, count_work=True)          # few reads, thousands of MEMs: partitions overflow, re-run
    compare(G, R, True); assert G["extensions"] == R["counters"]["E"]

def test_long_reads_with_indel_errors(capi):
    """BASELINE configs[4] as a parity case: 10 kbp reads with 5 % errors (substitutions + 1 bp indels): long backward /
    forward chains per lane, lanes of one warp finishing at very different times, many MEMs per read."""
    s = datasets.synth_index(ref_len=20000, n_hap=8, seed=21, encoded=True)
    g = capi.Index(s["ri"], s["tags0"], 0); o = orc.Index(s["ri"], s["tags0"], 0)
    reads, off = datasets.sample_reads(s["text"], 300, 10000, 0.04, seed=31, indel=0.01)
    G = g.find_mems(reads, off, 20, 1); R = o.find_mems(reads, off, 20, 1, count_work=True)
    compare(G, R, True)
    assert G["extensions"] == R["counters"]["E"] and G["n_mems"] > 300 * 50
    mixed = [reads[int(off[i]):int(off[i + 1])].tobytes()[: (10000 if i % 7 == 0 else 37 * (i % 11))] for i in range(120)]   # long and short lanes in one warp
    mr, mo = orc.pack_reads(mixed)
    compare(g.find_mems(mr, mo, 12, 2), o.find_mems(mr, mo, 12, 2), True)
    compare_query_tags(g.query_tags(mr, mo), o.query_tags(mr, mo), True)

@pytest.mark.parametrize("name", ["xy_legacy", "synth_enc", "synth_enc_big"])
def test_locate_batch(capi, name):
    """pgi_locate_batch = FastLocate::locate_encoded (toehold sample + phi) against the oracle; SA values
    row by row against decompressSA; sorted unique sequence ids; empty ranges; refusals."""
    prep = load_package().prep
    if name == "xy_legacy": ri = datasets.gold("xy.ri")
    elif name == "synth_enc": ri, _ = prep.rindex_from_rlbwt(datasets.synth_index()["rl_bwt"], encoded=True, with_samples=True)
    else: ri, _ = prep.rindex_from_rlbwt(datasets.synth_index(ref_len=20000, n_hap=8, seed=21)["rl_bwt"], encoded=True, with_samples=True)
    o = orc.Index(ri); g = capi.Index(ri); n = o.n
    assert g.info["has_locate"] == 1 and g.info["max_length"] == o.info[5]
    sa = o.decompress_sa()
    rng = np.random.default_rng(4); N = 20000
    a = rng.integers(0, n, N).astype(np.uint64); ln = np.where(rng.random(N) < 0.9, rng.integers(1, 50, N), rng.integers(50, 600, N)).astype(np.uint64)
    b = np.minimum(a + ln - 1, np.uint64(n - 1)); b[:50] = a[:50] - np.uint64(1) * (a[:50] > 0)      # empty (second < first) where a > 0
    a[50] = 0; b[50] = n - 1                                                                        # the whole BWT
    id_off, ids, row_off, rows = g.locate_batch(a, b)
    sizes = np.where(b >= a, b - a + 1, 0).astype(np.int64)
    assert np.array_equal(np.diff(row_off.astype(np.int64)), sizes)
    idx = np.repeat(a.astype(np.int64), sizes) + (np.arange(int(sizes.sum())) - np.repeat(row_off[:-1].astype(np.int64), sizes))
    assert np.array_equal(rows, sa[idx])                                                            # every row of every range
    k = 3000
    ooff, oids = o.locate_batch(a[:k], b[:k])
    assert np.array_equal(id_off[: k + 1], ooff) and np.array_equal(ids[: int(ooff[-1])], oids)
    for i in rng.integers(0, N, 200):
        assert np.array_equal(ids[id_off[i]:id_off[i + 1]], np.unique(sa[int(a[i]):int(b[i]) + 1] // np.uint64(o.info[5]))), i
    e = g.locate_batch(np.zeros(0, np.uint64), np.zeros(0, np.uint64)); assert len(e[1]) == 0 and len(e[3]) == 0
    with pytest.raises(capi.PgiError): g.locate_batch([0], [n])                                     # beyond the BWT
    with pytest.raises(capi.PgiError): capi.Index(datasets.synth_index()["ri"]).locate_batch([0], [1])   # file without samples
    # the FMD interval of a read and its located sequences agree with the text
    if name != "xy_legacy":
        text = datasets.synth_index(ref_len=20000, n_hap=8, seed=21)["text"] if name == "synth_enc_big" else datasets.synth_index()["text"]
        seqs = text.split(b"\n")[:-1]
        reads, off = datasets.sample_reads(text, 300, 25, 0.0, seed=8)
        q = g.query_tags(reads, off, text=False)["mems"]
        hit = q[:, 3] > 0
        lo, li, _, _ = g.locate_batch(q[hit, 2], q[hit, 2] + q[hit, 3] - np.uint64(1))
        for j, i in enumerate(np.nonzero(hit)[0][:100]):
            rd = reads[i * 25:(i + 1) * 25].tobytes()
            assert list(li[lo[j]:lo[j + 1]]) == [s for s in range(len(seqs)) if rd in seqs[s]], i

# ---- GPU-side construction: pgi_index_create_from_runs == pgi_index_create on the files prep writes for the same runs
def _runs_of_rlbwt(rl):
    sb, fb = np.frombuffer(rl[:16], dtype=np.uint64); rec = np.frombuffer(rl[16:], dtype=np.uint8).reshape(-1, int(sb + fb))
    ln = np.zeros(len(rec), dtype=np.uint64)
    for i in range(int(fb)): ln |= rec[:, 1 + i].astype(np.uint64) << np.uint64(8 * i)
    return rec[:, 0].copy(), ln

@pytest.mark.parametrize("shift", [None, "4", "6", "9", "12"])
def test_create_from_runs_equals_create_from_files(capi, shift, monkeypatch):
    if shift: monkeypatch.setenv("PGI_BSHIFT", shift)
    prep = load_package().prep
    s = datasets.synth_index(ref_len=20000, n_hap=8, seed=21, encoded=True)
    sym, ln = _runs_of_rlbwt(s["rl_bwt"])
    val = s["row_val"]; cut = np.nonzero(np.concatenate([[True], val[1:] != val[:-1]]))[0]
    tv = val[cut]; tl = np.diff(np.concatenate([cut, [len(val)]])).astype(np.uint64)
    tl_long = tl.copy(); tags_file, _ = prep.tags_from_runs(tv, tl_long, 0)
    a = capi.Index(s["ri"], tags_file, 0); b = capi.Index.from_runs(sym, ln, tv, tl_long)
    for k in ("n", "sigma", "is_encoded", "n_ref_blocks", "n_runs", "n_dev_blocks", "bucket_shift", "has_quirk_slot", "has_tags", "tag_layout",
              "tag_runs", "tag_items", "tag_item_width", "tag_universe", "n_crowded_buckets", "seed_k"):
        assert a.info[k] == b.info[k], k
    assert a.C == b.C and a.sym_map == b.sym_map
    if shift in ("9", "12"): assert b.info["n_crowded_buckets"] > 0
    pos = np.arange(0, a.n + 1, dtype=np.uint64)
    assert np.array_equal(a.rank_batch(pos), b.rank_batch(pos))
    reads, off = datasets.sample_reads(s["text"], 4000, 150, 0.01, seed=3)
    for ml, mo in ((20, 1), (4, 1)):
        compare(b.find_mems(reads, off, ml, mo), a.find_mems(reads, off, ml, mo), True)
    o = orc.Index(s["ri"], tags_file, 0)
    compare(b.find_mems(reads, off, 20, 1), o.find_mems(reads, off, 20, 1), True)
    st = np.arange(0, a.n - 8, 7, dtype=np.uint64); en = st + np.uint64(5)
    A = a.tag_query_batch(st, en); Bq = b.tag_query_batch(st, en)
    for k in ("n_runs", "tag_off", "tags"): assert np.array_equal(A[k], Bq[k]), k

def test_create_from_runs_long_tag_runs_and_errors(capi):
    """Tag runs longer than 511 are split (tag_arrays.cpp:942-957); bad symbols / zero lengths are refused."""
    prep = load_package().prep
    sym = np.frombuffer(b"\nACGT" * 40, dtype=np.uint8); ln = np.tile(np.array([1, 700, 30, 1200, 5], dtype=np.uint64), 40)
    n = int(ln.sum())
    tv = (np.arange(1, 41, dtype=np.uint64) << np.uint64(11)) | np.uint64(3); tl = np.full(40, n // 40, dtype=np.uint64); tl[-1] += n - int(tl.sum())
    ri, _ = prep.rindex_from_rlbwt(prep.rlbwt_from_runs(sym, ln), encoded=True, with_samples=False)
    tags_file, n_items = prep.tags_from_runs(tv, tl, 0)
    a = capi.Index(ri, tags_file, 0); b = capi.Index.from_runs(sym, ln, tv, tl)
    assert b.info["tag_items"] == n_items == a.info["tag_items"] and n_items > 40
    st = np.arange(0, n - 600, 13, dtype=np.uint64); en = st + np.uint64(590)
    A = a.tag_query_batch(st, en); Bq = b.tag_query_batch(st, en)
    for k in ("n_runs", "tag_off", "tags"): assert np.array_equal(A[k], Bq[k]), k
    pos = np.arange(0, n + 1, dtype=np.uint64); assert np.array_equal(a.rank_batch(pos), b.rank_batch(pos))
    with pytest.raises(capi.PgiError): capi.Index.from_runs(np.frombuffer(b"\nAXC", dtype=np.uint8), np.ones(4, dtype=np.uint64))
    with pytest.raises(capi.PgiError): capi.Index.from_runs(np.frombuffer(b"\nACG", dtype=np.uint8), np.array([1, 0, 2, 3], dtype=np.uint64))


def test_compact_result_equals_full_result(capi):
    """pgi_find_mems_compact (packed on the GPU, fewer bytes across PCIe) expands to exactly pgi_find_mems' arrays; fixed-length
    reads without an offset array and ragged reads with one."""
    for name in ("synth_enc_compact", "xy_legacy_bytecode", "two_contig_legacy"):
        c = get_case(name); g = capi.Index(c["ri"], c["tags"], c["layout"])
        reads, off = datasets.sample_reads(c["text"], 3000, 150, 0.01, seed=2)
        for ml in (20, 4):
            F = g.find_mems(reads, off, ml, 1, text=False); C = g.find_mems_compact(reads, 150, ml, 1)
            for k in ("n_mems", "n_tags", "extensions"): assert F[k] == C[k], k
            for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(F[k], C[k]), k
            assert C["wire_bytes"] < F["n_reads"] * 8 + F["n_mems"] * 48 + F["n_tags"] * 8
        rs = datasets.ragged_reads(c["text"], seed=3); buf, o2 = orc.pack_reads(rs)
        F = g.find_mems(buf, o2, 5, 1, text=False); C = g.find_mems_compact(buf, 0, 5, 1, read_off=o2)
        for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(F[k], C[k]), k

def test_bad_read_offsets_are_refused(capi):
    """offsets that are not ascending must not reach the kernel as wild addresses."""
    c = get_case("synth_enc_compact"); g = capi.Index(c["ri"], c["tags"], 0)
    reads, off = datasets.sample_reads(c["text"], 100, 150, 0.01, seed=2)
    bad = off.copy(); bad[50] = bad[40]
    with pytest.raises(capi.PgiError): g.find_mems(reads, bad, 20, 1, text=False)
    import torch                                                   # device-resident offsets: caught by the kernel, reported after the step
    d_reads = torch.cat([torch.from_numpy(reads), torch.zeros(16, dtype=torch.uint8)]).cuda(); d_bad = torch.from_numpy(bad.astype(np.int64)).cuda()
    b = capi.Batch(g); b.set_device_reads(d_reads.data_ptr(), d_bad.data_ptr(), 100, len(reads))
    with pytest.raises(capi.PgiError): b.run(20, 1)
    d_off = torch.from_numpy(off.astype(np.int64)).cuda(); b.set_device_reads(d_reads.data_ptr(), d_off.data_ptr(), 100, len(reads)); b.run(20, 1)   # the context survives
    compare(b.fetch(text=True), orc.Index(c["ri"], c["tags"], 0).find_mems(reads, off, 20, 1), True)

def test_clone_to_second_gpu(capi):
    """a replica on ANOTHER GPU (peer copy), queried there. Needs two devices."""
    if capi.lib().pgi_device_count() < 2: pytest.skip("one GPU visible")
    c = get_case("synth_enc_compact"); o = orc.Index(c["ri"], c["tags"], 0); g0 = capi.Index(c["ri"], c["tags"], 0, device=0)
    g1 = g0.clone_to(1)
    reads, off = datasets.sample_reads(c["text"], 3000, 150, 0.01, seed=4)
    R = o.find_mems(reads, off, 20, 1)
    compare(g1.find_mems(reads, off, 20, 1), R, True)
    compare(g0.find_mems(reads, off, 20, 1), R, True)
    s = datasets.synth_index(ref_len=20000, n_hap=8, seed=21, encoded=True)          # and one built from runs on device 1 directly
    sym, ln = _runs_of_rlbwt(s["rl_bwt"])
    b1 = capi.Index.from_runs(sym, ln, device=1)
    r2, o2 = datasets.sample_reads(s["text"], 1000, 150, 0.01, seed=5)
    compare(b1.find_mems(r2, o2, 20, 1), orc.Index(s["ri"]).find_mems(r2, o2, 20, 1, with_tags=False), False)

def test_async_steps_overlap_and_compact_in_step(capi):
    """pgi_batch_run_async / pgi_batch_wait: three batches in flight on one GPU from one host thread return what
    pgi_batch_run returns for each of them (incl. a first step whose buffers must grow: the re-run happens inside wait), and
    pgi_find_mems_compact (result packed inside the step, steps of concurrent callers chained on the GPU) expands to the
    arrays of pgi_find_mems, for every PGI_CHAIN mode."""
    import os, threading
    s = datasets.synth_index(ref_len=20000, n_hap=8, seed=21, encoded=True)
    g = capi.Index(s["ri"], s["tags0"], 0)
    sets = [datasets.sample_reads(s["text"], 120000 + 7000 * i, 150, 0.02, seed=30 + i) for i in range(3)]
    want = []
    for reads, off in sets:
        b = capi.Batch(g); b.upload(reads.ctypes.data, off.ctypes.data, len(off) - 1); b.run(20, 1); want.append(b.fetch()); b.close()
    bs = [capi.Batch(g) for _ in sets]
    for rep in range(2):                                       # second round: the batches' buffers are sized, no re-run
        for b, (reads, off) in zip(bs, sets): b.upload(reads.ctypes.data, off.ctypes.data, len(off) - 1); b.run_async(20, 1)
        for b, W in zip(bs, want):
            b.wait(); S = b.fetch()
            for k in ("n_mems", "n_tags", "extensions"): assert S[k] == W[k], k
            for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(S[k], W[k]), k
    with pytest.raises(capi.PgiError): bs[0].wait()             # nothing enqueued
    old = os.environ.get("PGI_CHAIN")
    try:
        for mode in ("1", "0", "2"):
            os.environ["PGI_CHAIN"] = mode
            outs = [None] * 6
            def call(i):
                reads, off = sets[i % 3]
                outs[i] = g.find_mems_compact(reads, 150, 20, 1, read_off=(off if i == 1 else None)) if i < 3 else g.find_mems(reads, off, 20, 1, text=False)
            ths = [threading.Thread(target=call, args=(i,)) for i in range(6)]
            [t.start() for t in ths]; [t.join() for t in ths]
            for i, o_ in enumerate(outs):
                for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(o_[k], want[i % 3][k]), (mode, i, k)
    finally:
        if old is None: os.environ.pop("PGI_CHAIN", None)
        else: os.environ["PGI_CHAIN"] = old

@pytest.mark.parametrize("variant", ["44", "24", "14", "04"])
def test_find_mems_forced_kernel_variants(capi, variant):
    """Every block-fetch mode of fmd_mem_kernel (PGI_K1; the default is chosen by index size: mode 1 for an L2-resident
    index, mode 4 otherwise) returns the oracle's MEMs, tags and extension count on the synthetic and the shipped indexes."""
    import os
    old = os.environ.get("PGI_K1"); os.environ["PGI_K1"] = variant
    try:
        for name in ("synth_enc_compact", "xy_legacy_bytecode", "x_encoded_sigma5"):
            c = get_case(name)
            o = orc.Index(c["ri"], c["tags"], c["layout"]); g = capi.Index(c["ri"], c["tags"], c["layout"])
            reads, off = datasets.sample_reads(c["text"], 3000, 150, 0.02, seed=17)
            for ml, mo in ((20, 1), (5, 2)):
                G = g.find_mems(reads, off, ml, mo); R = o.find_mems(reads, off, ml, mo, count_work=True)
                compare(G, R, c["tags"] is not None)
                assert G["extensions"] == R["counters"]["E"]
    finally:
        if old is None: os.environ.pop("PGI_K1", None)
        else: os.environ["PGI_K1"] = old
