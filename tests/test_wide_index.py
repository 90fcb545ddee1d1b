"""Indexes with n > 2^32: the 40-bit position paths (block headers' high bytes pgi_core.cuh pos_rank, seed_pack, 64-bit
interval arithmetic, tag universes beyond 2^32), which no real text small enough for a test reaches. Rank / extension /
MEM-finding parity needs a well-formed .ri, not a real BWT: the run-length BWT is FABRICATED (random symbols, run lengths
up to 2^20, a few hundred end markers, some N runs), pushed through the same writers (prep) and compared with the oracle."""
import numpy as np, pytest
import orc, datasets
from conftest import load_package

def fabricate_runs(n_target, seed, max_len=1 << 20):
    rng = np.random.default_rng(seed)
    r = int(n_target / (max_len / 2)) + 64
    ln = rng.integers(1, max_len, r).astype(np.uint64)
    sym = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, r)].copy()
    for i in range(1, r):                                   # maximal runs: no two neighbours alike
        if sym[i] == sym[i - 1]: sym[i] = b"ACGT"[(b"ACGT".index(bytes([sym[i]])) + 1) % 4]
    nl = rng.choice(r, 300, replace=False); sym[nl] = 10; ln[nl] = 1          # every end marker is its own run
    nn = rng.choice(np.setdiff1d(np.arange(r), np.concatenate([nl, nl + 1, nl - 1])), 40, replace=False)
    for i in nn:
        if sym[i - 1] != 78 and (i + 1 >= r or sym[i + 1] != 78): sym[i] = 78
    keep = np.ones(r, dtype=bool); keep[1:] = sym[1:] != sym[:-1]; keep |= sym == 10
    return sym[keep], ln[keep]

def random_queries(n, seed, N):
    rng = np.random.default_rng(seed)
    iv = np.zeros((N, 3), dtype=np.uint64)
    k = rng.integers(0, n, N); sz = np.minimum(rng.integers(0, 1 << 22, N), n - k)
    iv[:, 0] = k; iv[:, 1] = rng.integers(0, n, N); iv[:, 2] = sz
    iv[:200] = [0, 0, n]
    iv[200:400, 2] = np.minimum(rng.integers(0, n, 200), n - iv[200:400, 0])          # very long intervals
    sym = rng.choice(np.frombuffer(b"ACGTN\n\0acgtnX", dtype=np.uint8), N); d = rng.integers(0, 2, N).astype(np.uint8)
    return iv, sym, d

def random_reads(n_reads, L, seed):
    rng = np.random.default_rng(seed)
    r = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, n_reads * L)].copy()
    r[rng.random(n_reads * L) < 0.002] = 78
    return r, np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(L)

def test_hostsim_33_bit_index():
    """CPU: the product's device arithmetic (hostsim) on n = 2^32 + 2^29 against the oracle."""
    import sim
    prep = load_package().prep
    sym, ln = fabricate_runs((1 << 32) + (1 << 29), seed=1)
    n = int(ln.sum()); assert n > 1 << 32
    ri, info = prep.rindex_from_rlbwt(prep.rlbwt_from_runs(sym, ln), encoded=True, with_samples=False)
    o = orc.Index(ri); g = sim.Index(ri)
    assert g.info["n"] == n == o.n and g.info["sigma"] == 6
    iv, s, d = random_queries(n, 2, 30000)
    want = o.extend_batch(iv, s, d)
    assert np.array_equal(g.extend_batch(iv, s, d), want)
    assert (want[:, 0] >= (1 << 32)).sum() > 300 and (want[:, 1] >= (1 << 32)).sum() > 300      # intervals beyond 2^32 do occur
    assert g.set_seed_k(8) == 8                                                                # 40-bit seed entries
    reads, off = random_reads(60, 150, 3)
    R = o.find_mems(reads, off, 9, 1, with_tags=False, count_work=True, threads=1)
    tot = 0; big = 0
    for i in range(60):
        mems, ne = g.find_mems(reads[int(off[i]):int(off[i + 1])].tobytes(), 9, 1); tot += ne
        assert np.array_equal(mems, R["mems"][int(R["mem_off"][i]):int(R["mem_off"][i + 1])]), i
        big += int((mems[:, 2] >= (1 << 32)).sum())
    assert tot == R["counters"]["E"] and big > 10

@pytest.fixture(scope="module")
def capi():
    c = load_package().capi
    assert c.lib().pgi_device_count() > 0, "no CUDA device visible: the product has no CPU path"
    return c

@pytest.mark.gpu
def test_gpu_37_bit_index_from_runs(capi):
    """n = 1.2 * 2^36 (config 4's BWT is 2^36.2 long): built on the GPU from the runs, checked against the oracle on the .ri
    the writers emit for the same runs. Every byte of the 40-bit fields that config 4 uses is exercised."""
    prep = load_package().prep
    sym, ln = fabricate_runs(int(1.2 * (1 << 36)), seed=5)
    n = int(ln.sum()); assert n > 1 << 36
    ri, info = prep.rindex_from_rlbwt(prep.rlbwt_from_runs(sym, ln), encoded=True, with_samples=False)
    o = orc.Index(ri); g = capi.Index.from_runs(sym, ln)
    assert g.info["n"] == n and g.info["seed_k"] == 13 and g.info["sigma"] == 6 and g.C == o.C
    iv, s, d = random_queries(n, 6, 50000)
    want = o.extend_batch(iv, s, d)
    assert np.array_equal(g.extend_batch(iv, s, d), want)
    assert (want[:, 0] >= (1 << 35)).sum() > 1000 and (want[:, 1] >= (1 << 35)).sum() > 1000
    pos = np.concatenate([np.random.default_rng(7).integers(0, n + 1, 20000), [0, n, n - 1, 1 << 32, (1 << 36) - 1, 1 << 36]]).astype(np.uint64)
    assert np.array_equal(g.rank_batch(pos), o.rank_batch(pos))
    reads, off = random_reads(20000, 150, 8)
    for ml in (14, 20):
        R = o.find_mems(reads, off, ml, 1, with_tags=False, count_work=True); G = g.find_mems(reads, off, ml, 1, text=False)
        assert G["n_mems"] == R["n_mems"] and np.array_equal(G["mem_off"], R["mem_off"]) and np.array_equal(G["mems"], R["mems"])
        assert G["extensions"] == R["counters"]["E"] and G["seed_lookups"] > 0
        if ml == 14: assert (R["mems"][:, 2] >= (1 << 35)).sum() > 100                            # MEM intervals far beyond 2^32
    Q = g.query_tags(reads[: 150 * 2000].reshape(-1, 150)[:, :18].reshape(-1).copy(), np.arange(2001, dtype=np.uint64) * np.uint64(18), text=False)
    QR = o.query_tags(reads[: 150 * 2000].reshape(-1, 150)[:, :18].reshape(-1).copy(), np.arange(2001, dtype=np.uint64) * np.uint64(18), text=False, with_tags=False)
    assert np.array_equal(Q["mems"], QR["mems"])

@pytest.mark.gpu
def test_gpu_tag_universe_beyond_2_32(capi):
    """Tag arrays over 2^33 BWT positions (runs up to 2^20 long, split at 511): rank / select / decode beyond 2^32."""
    prep = load_package().prep
    rng = np.random.default_rng(9)
    tl = rng.integers(1, 1 << 20, 18000).astype(np.uint64); tv = (rng.integers(1, 1 << 24, len(tl)).astype(np.uint64) << np.uint64(11)) | rng.integers(0, 1024, len(tl)).astype(np.uint64)
    total = int(tl.sum()); assert total > 1 << 33
    tags, n_items = prep.tags_from_runs(tv, tl, 0)
    o = orc.Index(datasets.gold("xy.ri"), tags, 0); g = capi.Index(None, tags, 0)
    assert g.info["tag_universe"] == total + 1 and g.info["tag_items"] == n_items
    st = rng.integers(0, total - 5000, 3000).astype(np.uint64); st[:1500] += np.uint64(1 << 32); st = np.minimum(st, np.uint64(total - 5000))
    en = st + rng.integers(0, 4000, 3000).astype(np.uint64)
    G = g.tag_query_batch(st, en)
    for i in range(len(st)):
        nr, tvv = o.tag_query(int(st[i]), int(en[i]))
        a, b = G["tag_off"][i], G["tag_off"][i + 1]
        assert G["n_runs"][i] == nr and np.array_equal(G["tags"][a:b], tvv), i

@pytest.mark.gpu
def test_gpu_nominal_config3_properties(capi):
    """BASELINE config 3 at NOMINAL size (64 haplotypes x 50 Mbp + reverse complements, n = 6.4e9, built from the graph's
    structure on the GPU): size-independent properties of a 400 000-read batch, and every sampled MEM's interval recomputed
    through an independent path -- character by character with pgi_extend_batch (lane-private block loads, no seed table, no
    cooperative fetch) -- must equal what the MEM kernel reported; each MEM is right-maximal; a second run is identical."""
    import torch
    pkg = load_package(); pg = pkg.pangraph
    d = pg.build(50_000_000, 64, seed=3, device=0, host=False)
    assert d["n"] > 6_300_000_000
    r = d["dev_runs"]; g = capi.Index.from_runs(r["sym"], r["len"], r["tag_val"], r["tag_len"], device=0)
    n = g.info["n"]; assert n == d["n"] and g.info["has_tags"] == 1
    N, L, ML = 400_000, 150, 20
    reads = pg.sample_reads(d["strands"], N, L, 0.01, seed=4).reshape(-1).cpu().numpy(); off = np.arange(N + 1, dtype=np.uint64) * np.uint64(L)
    del d, r; torch.cuda.empty_cache()
    G = g.find_mems(reads, off, ML, 1, text=False)
    G2 = g.find_mems(reads, off, ML, 1, text=False)
    for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(G[k], G2[k]), k           # idempotent
    m = G["mems"]; per_read = np.diff(G["mem_off"].astype(np.int64)); rid = np.repeat(np.arange(N), per_read)
    assert G["n_mems"] > N and (m[:, 1] - m[:, 0] >= ML).all() and (m[:, 1] <= L).all() and (m[:, 3].astype(np.int64) > 0).all()
    assert (m[:, 2] + m[:, 3] <= n).all() and (m[:, 2] >= (1 << 32)).sum() > 1000
    same = rid[1:] == rid[:-1]; assert (m[1:, 0][same] > m[:-1, 0][same]).all()                                  # discovery order
    t0 = G["tag_off"].astype(np.int64); assert (np.diff(t0) >= 1).all() and (np.diff(t0) <= G["n_runs"].astype(np.int64)).all()
    tg = G["tags"]; inner = np.ones(len(tg), dtype=bool); inner[t0[:-1]] = False
    assert (tg[1:][inner[1:]] > tg[:-1][inner[1:]]).all()                                                        # sorted unique per MEM
    rng = np.random.default_rng(1); pick = rng.choice(G["n_mems"], 4000, replace=False)
    iv = np.zeros((len(pick), 3), dtype=np.uint64); iv[:, 2] = n
    st = m[pick, 0].astype(np.int64); en = m[pick, 1].astype(np.int64); base = rid[pick].astype(np.int64) * L
    alive = np.ones(len(pick), dtype=bool)
    for step in range(int((en - st).max())):
        pos = en - 1 - step; act = pos >= st
        if not act.any(): break
        sym = np.where(act, reads[base + np.maximum(pos, 0)], ord("A")).astype(np.uint8)
        out = g.extend_batch(iv, sym, np.zeros(len(pick), dtype=np.uint8))
        iv[act] = out[act]
    assert np.array_equal(iv[:, 0], m[pick, 2]) and np.array_equal(iv[:, 2], m[pick, 3])                          # same interval, independent path
    more = en < L                                                                                                  # right-maximal: one more character fails
    rev = np.zeros((len(pick), 3), dtype=np.uint64)                                                               # forward extension needs the reverse interval: rebuild it by a forward walk
    rev[:, 2] = n
    for step in range(int((en - st).max())):
        pos = st + step; act = pos < en
        sym = np.where(act, reads[base + np.minimum(pos, L - 1)], ord("A")).astype(np.uint8)
        out = g.extend_batch(rev, sym, np.ones(len(pick), dtype=np.uint8))
        rev[act] = out[act]
    assert np.array_equal(rev[:, 0], m[pick, 2]) and np.array_equal(rev[:, 2], m[pick, 3])                        # forward walk reaches the same interval
    nxt = np.where(more, reads[base + np.minimum(en, L - 1)], ord("A")).astype(np.uint8)
    out = g.extend_batch(rev, nxt, np.ones(len(pick), dtype=np.uint8))
    assert (out[more, 2] == 0).all()
