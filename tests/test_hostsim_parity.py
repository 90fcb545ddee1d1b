"""CPU: the PRODUCT's device arithmetic (csrc/pgi_core.cuh: block locate, packed run scan, FMD extension, MEM lane
state machine, tag plan / item decoders) and host flattener (csrc/pgi_flatten.hpp), compiled as plain C++ by
tests/hostsim (test-only), against the oracle. This is the GPU-less safety net for kernel changes; the same
comparisons run on the GPU through the C ABI in test_gpu_parity.py."""
import numpy as np, pytest
import orc, sim, datasets

def case(name):
    if name == "xy_legacy_bytecode":
        return dict(ri=datasets.gold("xy.ri"), tags=datasets.gold("xy_bidirectional_compressed.tags"), layout=1, text=datasets.gold("contigs_xy"))
    if name == "two_contig_legacy":
        return dict(ri=datasets.gold("two_contig_xy.ri"), tags=None, layout=0, text=datasets.gold("contigs_XY.txt"))
    enc = name != "synth_leg_compact"
    s = datasets.synth_index(encoded=enc)
    if name == "synth_enc_bytecode":
        return dict(ri=s["ri"], tags=s["tags1"], layout=1, text=s["text"])
    return dict(ri=s["ri"], tags=s["tags0"], layout=0, text=s["text"])

NAMES = ["xy_legacy_bytecode", "two_contig_legacy", "synth_enc_compact", "synth_leg_compact", "synth_enc_bytecode"]

@pytest.mark.parametrize("name", NAMES)
def test_flatten_facts(name):
    c = case(name); o = orc.Index(c["ri"], c["tags"], c["layout"]); g = sim.Index(c["ri"], c["tags"], c["layout"])
    assert g.info["n"] == o.n and g.info["sigma"] == o.sigma and g.info["encoded"] == o.info[2]
    assert g.info["n_ref_blocks"] == o.info[4]
    assert g.info["has_quirk"] == (1 if name in ("xy_legacy_bytecode", "two_contig_legacy") else 0)   # legacy + sigma 5
    if c["tags"]:
        assert g.info["tag_m"] == o.info[12]

@pytest.mark.parametrize("name", NAMES)
def test_extend_matches_oracle(name):
    c = case(name); o = orc.Index(c["ri"]); g = sim.Index(c["ri"])
    rng = np.random.default_rng(3); N = 40000; n = o.n
    iv = np.zeros((N, 3), dtype=np.uint64)
    k = rng.integers(0, n, N); sz = np.minimum(rng.integers(0, 400, N), n - k)
    iv[:, 0] = k; iv[:, 1] = rng.integers(0, n, N); iv[:, 2] = sz
    iv[:300] = [0, 0, n]
    iv[300:400, 0] = rng.integers(0, 2 * n, 100)             # positions beyond n (defined UB: read as n)
    iv[400:500, 2] = rng.integers(0, n, 100)                 # long intervals
    iv[400:500, 0] = np.minimum(iv[400:500, 0], n - iv[400:500, 2])
    sym = rng.choice(np.frombuffer(b"ACGTN\n\0acgtnX\xff", dtype=np.uint8), N); d = rng.integers(0, 2, N).astype(np.uint8)
    assert np.array_equal(g.extend_batch(iv, sym, d), o.extend_batch(iv, sym, d))

@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("params", [(20, 1), (5, 1), (1, 1), (10, 3)])
def test_lane_machine_matches_find_all_mems(name, params):
    c = case(name); ml, mo = params
    o = orc.Index(c["ri"]); g = sim.Index(c["ri"])
    nreads = 300 if ml >= 10 else 60
    reads, off = datasets.sample_reads(c["text"], nreads, 150, 0.02, seed=ml + 3 * mo)
    R = o.find_mems(reads, off, ml, mo, with_tags=False, count_work=True, threads=1)
    tot_ext = 0
    for i in range(nreads):
        rd = reads[int(off[i]):int(off[i + 1])].tobytes()
        mems, ne = g.find_mems(rd, ml, mo); tot_ext += ne
        assert np.array_equal(mems, R["mems"][int(R["mem_off"][i]):int(R["mem_off"][i + 1])]), i
    assert tot_ext == R["counters"]["E"]
    for rd in datasets.ragged_reads(c["text"], seed=2, n=40):
        buf, o2 = orc.pack_reads([rd])
        want = o.find_mems(buf, o2, ml, mo, with_tags=False)["mems"]
        got, _ = g.find_mems(rd, ml, mo)
        assert np.array_equal(got, want), rd

@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("K", [1, 3, 5, 6, 8])
def test_seed_table_leaves_find_all_mems_unchanged(name, K):
    """The seed table (pgi_core.cuh: one lookup for the first K backward extensions of a walk from the full interval) must
    not change a single MEM nor the extension count, for every K, min_len below / equal / above K, min_occ > 1 (entries
    that fail the occurrence test fall back to single steps), reads with other bytes, ragged and empty reads."""
    c = case(name); o = orc.Index(c["ri"]); g = sim.Index(c["ri"])
    k = g.set_seed_k(K)
    if g.info["has_quirk"]:                              # quirk slot: no table (sizes are not monotone there)
        assert k == 0; return
    assert k == K
    for ml, mo in [(20, 1), (K, 1), (max(1, K - 1), 1), (K + 1, 2), (12, 40)]:
        nreads = 120
        reads, off = datasets.sample_reads(c["text"], nreads, 150, 0.03, seed=K + ml + mo)
        R = o.find_mems(reads, off, ml, mo, with_tags=False, count_work=True, threads=1)
        tot_ext = 0
        for i in range(nreads):
            rd = reads[int(off[i]):int(off[i + 1])].tobytes()
            mems, ne = g.find_mems(rd, ml, mo); tot_ext += ne
            assert np.array_equal(mems, R["mems"][int(R["mem_off"][i]):int(R["mem_off"][i + 1])]), (ml, mo, i)
        assert tot_ext == R["counters"]["E"], (ml, mo)
        for rd in datasets.ragged_reads(c["text"], seed=5, n=40):
            buf, o2 = orc.pack_reads([rd])
            want = o.find_mems(buf, o2, ml, mo, with_tags=False)["mems"]
            got, _ = g.find_mems(rd, ml, mo)
            assert np.array_equal(got, want), (ml, mo, rd)
    g.set_seed_k(0)

@pytest.mark.parametrize("name", ["xy_legacy_bytecode", "synth_enc_compact", "synth_enc_bytecode"])
def test_tag_plan_matches_oracle(name):
    c = case(name); o = orc.Index(c["ri"], c["tags"], c["layout"]); g = sim.Index(c["ri"], c["tags"], c["layout"])
    rng = np.random.default_rng(11); n = o.n
    for _ in range(3000):
        a = int(rng.integers(0, n)); b = min(n + 30, a + int(rng.integers(0, 50 if rng.random() < 0.9 else 2500)))
        nr, tv = o.tag_query(a, b); nr2, tv2 = g.tag_query(a, b)
        assert nr == nr2 and np.array_equal(tv, tv2), (a, b)
    for a in range(0, min(n, 1500)):                          # every alignment of the first % 10 rule
        nr, tv = o.tag_query(a, a + 2); nr2, tv2 = g.tag_query(a, a + 2)
        assert nr == nr2 and np.array_equal(tv, tv2), a

@pytest.mark.parametrize("name", NAMES)
def test_count_matches_oracle(name):
    """query_tags path: whole-read LF chain (count descriptors) against the oracle's count_encoded / count restatement."""
    c = case(name); o = orc.Index(c["ri"]); g = sim.Index(c["ri"])
    rng = np.random.default_rng(5)
    lines = [l for l in c["text"].split(b"\n") if len(l) > 100]
    rs = [b"", b"A", b"T", b"N", b"\n", b"a", b"ACGT", bytes([0, 200]), b"GATTACA"]
    for _ in range(400):
        L = lines[int(rng.integers(len(lines)))]; ln = int(rng.integers(1, 60)); st = int(rng.integers(0, len(L) - ln))
        r = bytearray(L[st:st + ln])
        if rng.random() < 0.3: r[int(rng.integers(len(r)))] = b"ACGTN\nx"[int(rng.integers(7))]
        rs.append(bytes(r))
    buf, off = orc.pack_reads(rs)
    R = o.query_tags(buf, off, with_tags=False, threads=1)
    hits = 0
    for i, rd in enumerate(rs):
        got = g.count(rd)
        assert got is not None
        assert got == (int(R["mems"][i, 2]), int(R["mems"][i, 3])), (i, rd)
        hits += got[1] > 0
    assert hits > 100

@pytest.mark.parametrize("name", ["xy_legacy", "two_contig_legacy", "synth_enc", "synth_leg"])
def test_locate_matches_oracle(name):
    """Locate (toehold sample of the run + phi steps) over the flattened samples / last / last_to_run against the oracle's
    locate_encoded / decompressSA; refused on an encoded file without N and on files without samples."""
    from conftest import load_package
    prep = load_package().prep
    if name == "xy_legacy": ri = datasets.gold("xy.ri")
    elif name == "two_contig_legacy": ri = datasets.gold("two_contig_xy.ri")
    else: ri, _ = prep.rindex_from_rlbwt(datasets.synth_index()["rl_bwt"], encoded=name == "synth_enc", with_samples=True)
    o = orc.Index(ri); g = sim.Index(ri); sa = o.decompress_sa()
    assert np.array_equal(g.locate(0, o.n - 1), sa)
    rng = np.random.default_rng(9)
    for _ in range(500):
        a = int(rng.integers(0, o.n)); b = min(o.n - 1, a + int(rng.integers(0, 60)))
        _, _, rows = o.locate_batch([a], [b], want_sa=True)
        assert np.array_equal(g.locate(a, b), rows) and np.array_equal(rows, sa[a:b + 1]), (a, b)
    assert sim.Index(datasets.synth_index()["ri"]).locate(0, 3) is None                 # built without samples
    s5 = datasets.synth_index(ref_len=2000, n_hap=2, seed=3, encoded=True, n_N=0)
    ri5, _ = prep.rindex_from_rlbwt(s5["rl_bwt"], encoded=True, with_samples=True)
    assert sim.Index(ri5).locate(0, 3) is None                                           # encoded, sigma 5


@pytest.mark.parametrize("shift", ["4", "8", "10", "12"])
def test_forced_bucket_shift(shift, monkeypatch):
    """Bucket sizes from 16 to 4096 positions (PGI_BSHIFT): with the large ones nearly every bucket of these small indexes is
    crowded (directory block + sub-bucket chunks, pgi_layout.h); extension and MEM finding must not notice."""
    monkeypatch.setenv("PGI_BSHIFT", shift)
    for name in ("xy_legacy_bytecode", "synth_enc_compact"):
        c = case(name); o = orc.Index(c["ri"]); g = sim.Index(c["ri"])
        assert g.info["bshift"] == int(shift)
        if int(shift) >= 8: assert g.info["n_crowded"] > 0
        rng = np.random.default_rng(5); N = 20000; n = o.n
        iv = np.zeros((N, 3), dtype=np.uint64)
        k = rng.integers(0, n, N); sz = np.minimum(rng.integers(0, 600, N), n - k)
        iv[:, 0] = k; iv[:, 1] = rng.integers(0, n, N); iv[:, 2] = sz; iv[:200] = [0, 0, n]
        sym = rng.choice(np.frombuffer(b"ACGTN\n\0", dtype=np.uint8), N); d = rng.integers(0, 2, N).astype(np.uint8)
        assert np.array_equal(g.extend_batch(iv, sym, d), o.extend_batch(iv, sym, d))
        pos_iv = np.zeros((n + 1, 3), dtype=np.uint64); pos_iv[:, 0] = np.arange(n + 1); pos_iv[:, 2] = 0        # every position as k (size 0 -> k + s = k)
        s2 = np.full(n + 1, ord("A"), dtype=np.uint8); d2 = np.zeros(n + 1, dtype=np.uint8)
        assert np.array_equal(g.extend_batch(pos_iv, s2, d2), o.extend_batch(pos_iv, s2, d2))
        reads, off = datasets.sample_reads(c["text"], 60, 150, 0.02, seed=4)
        R = o.find_mems(reads, off, 12, 1, with_tags=False, threads=1)
        for i in range(60):
            mems, ne = g.find_mems(reads[int(off[i]):int(off[i + 1])].tobytes(), 12, 1)
            assert np.array_equal(mems, R["mems"][int(R["mem_off"][i]):int(R["mem_off"][i + 1])]), i

def test_header_limb_sums_at_the_limits_of_the_layout():
    """pos_rank's header arithmetic (counts stored as 16-bit limbs, weighted sums as dp2a / dp4a against the descriptor's byte
    tables) against plain integer arithmetic: random and extreme counts up to 2^40 - 1 (A, C, G, T) / 2^32 - 1 (NL, N), every
    weight 0..7, every position of the main flag."""
    import ctypes, sim
    L = sim.lib(); L.sim_header_sums.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p]
    rng = np.random.default_rng(5)
    lim = [(1 << 32) - 1, (1 << 40) - 1, (1 << 40) - 1, (1 << 40) - 1, (1 << 32) - 1, (1 << 40) - 1]     # code order: NL A C G N T
    out = np.zeros(2, dtype=np.uint64)
    cases = [np.array(lim, dtype=np.uint64), np.zeros(6, dtype=np.uint64), np.array([1, 1 << 16, 1 << 32, (1 << 32) - 1, 1 << 31, (1 << 39) + 12345], dtype=np.uint64)]
    cases += [np.array([int(rng.integers(0, l + 1)) for l in lim], dtype=np.uint64) for _ in range(400)]
    for k, cnt in enumerate(cases):
        wts = [7] * 6 if k == 0 else [int(x) for x in rng.integers(0, 8, 6)]
        main = int(rng.integers(0, 7))                                   # 6: no main symbol
        tbl = 0
        for c in range(6): tbl |= (wts[c] | (8 if c == main else 0)) << (4 * c)
        L.sim_header_sums(cnt.ctypes.data, tbl, out.ctypes.data)
        W = sum(int(cnt[c]) * wts[c] for c in range(6)); M = int(cnt[main]) if main < 6 else 0
        assert int(out[0]) == W % (1 << 64) and int(out[1]) == M, (k, cnt, wts, main)
