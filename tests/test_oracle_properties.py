"""CPU: properties that pin the oracle independently of any index structure, on sigma = 6 bidirectional texts
(the regime where the reference's rank vector is sane):
  * the two FMD properties of the reference's tests/test_rindex.cpp:288-486 (interval size = occurrence count,
    k-mer / reverse-complement intervals are mirror images);
  * legacy and encoded files of the same BWT answer identically;
  * find_all_mems equals a brute-force restatement that counts occurrences in the plain text;
  * tag queries equal a brute-force lookup over per-row values (away from the first % 10 == 0 quirk)."""
import numpy as np, pytest
import orc, datasets

def occurrences(text: bytes, pat: bytes) -> int:
    c = 0; i = text.find(pat)
    while i >= 0:
        c += 1; i = text.find(pat, i + 1)
    return c

@pytest.fixture(scope="module")
def syn():
    s = datasets.synth_index(ref_len=1500, n_hap=3, seed=11, encoded=True)
    l = datasets.synth_index(ref_len=1500, n_hap=3, seed=11, encoded=False)
    return s, l

def test_sigma6_and_sizes(syn):
    s, l = syn
    o = orc.Index(s["ri"]); assert o.sigma == 6 and o.info[2] == 1 and o.info[3] == 1
    ol = orc.Index(l["ri"]); assert ol.sigma == 6 and ol.info[2] == 0
    assert o.n == len(s["text"]) == ol.n

def test_interval_size_is_occurrence_count_and_mirror(syn):
    s, _ = syn
    o = orc.Index(s["ri"]); text = s["text"]; rng = np.random.default_rng(42)
    line = s["seqs"][0]
    for _ in range(150):
        k = int(rng.integers(1, 14)); st = int(rng.integers(0, len(line) - k)); pat = line[st:st + k]
        f, r, sz = o.search(pat)
        assert sz == occurrences(text, pat), pat
        f2, r2, sz2 = o.search(datasets.revcomp(pat))
        if b"N" not in pat:
            assert (f2, r2, sz2) == (r, f, sz), pat                  # FMD mirror property
    # extending the full interval with '\0' behaves like '\n' (sym_map aliasing)
    full = np.array([[0, 0, o.n]], dtype=np.uint64)
    a = o.extend_batch(full, [0], [0]); b = o.extend_batch(full, [10], [0])
    assert np.array_equal(a, b) and a[0, 2] == len(s["seqs"])

def test_backward_extension_matches_rank_difference(syn):
    """BackwardExtensionMatchesLF (reference tests/test_rindex.cpp:288-373): size / forward of backward_extend
    follow from the rank vector: forward = C[a] + R(k)[a], size = R(k+s)[a] - R(k)[a]."""
    s, _ = syn
    o = orc.Index(s["ri"]); pat = s["seqs"][2][100:172]; assert len(pat) == 72
    iv = np.array([[0, 0, o.n]], dtype=np.uint64)
    for ch in reversed(pat):
        k, sz = int(iv[0, 0]), int(iv[0, 2])
        R = o.rank_batch([k, k + sz]); a = o.sym_map[ch]
        iv = o.extend_batch(iv, [ch], [0])
        assert int(iv[0, 0]) == o.C[a] + int(R[0, a]) and int(iv[0, 2]) == int(R[1, a]) - int(R[0, a])

def test_legacy_equals_encoded(syn):
    s, l = syn
    a = orc.Index(s["ri"], s["tags0"], 0); b = orc.Index(l["ri"], l["tags0"], 0)
    pos = np.arange(0, a.n + 1, dtype=np.uint64)
    assert np.array_equal(a.rank_batch(pos), b.rank_batch(pos))
    reads, off = datasets.sample_reads(s["text"], 300, 100, 0.03, seed=3)
    for ml, mo in [(12, 1), (4, 2)]:
        A = a.find_mems(reads, off, ml, mo); B = b.find_mems(reads, off, ml, mo)
        assert A["text"] == B["text"] and A["n_mems"] > 0

def brute_find_all_mems(text: bytes, read: bytes, min_len: int, min_occ: int):
    """include/pangenome_index/algorithm.hpp:653-757 with every FMD interval size replaced by a plain-text
    occurrence count (valid for sigma = 6 bidirectional texts). '\\0' / absent bytes alias the endmarker."""
    present = set(text)
    def canon(b):                       # bytes absent from the text alias '\n' (sym_map == 0)
        return bytes(c if c in present else 10 for c in b)
    def cnt(p): return occurrences(text, canon(p))
    L = len(read); pat = read + b"\0"; out = []; x = 0
    while x < L:
        if L - x < min_len: break
        j = x + min_len - 1; fail = None
        while j >= x:
            if cnt(pat[j:x + min_len]) < max(min_occ, 1): fail = j + 1; break
            j -= 1
        if fail is not None: x = fail; continue
        e = x + min_len
        while e < L and cnt(pat[x:e + 1]) >= max(min_occ, 1): e += 1
        out.append((x, e, cnt(pat[x:e])))
        j = e; nxt = None
        while j > x:
            if cnt(pat[j:e + 1]) < max(min_occ, 1): nxt = j + 1; break
            j -= 1
        x = nxt if nxt is not None else x + 1
    return out

@pytest.mark.parametrize("ml,mo", [(8, 1), (5, 2), (15, 1), (3, 4)])
def test_mems_equal_brute_force(syn, ml, mo):
    s, _ = syn
    o = orc.Index(s["ri"]); text = s["text"]
    reads, off = datasets.sample_reads(text, 25, 60, 0.05, seed=ml + mo)
    R = o.find_mems(reads, off, ml, mo, with_tags=False)
    for i in range(25):
        rd = reads[int(off[i]):int(off[i + 1])].tobytes()
        got = [(int(m[0]), int(m[1]), int(m[3])) for m in R["mems"][int(R["mem_off"][i]):int(R["mem_off"][i + 1])]]
        assert got == brute_find_all_mems(text, rd, ml, mo), (i, rd)

def tag_runs(val):
    """(start row, value) of every tag run; prep splits runs longer than 511 (tag_arrays.cpp:942-957)."""
    runs = []; prev = None; ln = 0
    for p, v in enumerate(val):
        v = int(v)
        if v != prev or ln == 511: runs.append((p, v)); prev = v; ln = 0
        ln += 1
    return runs

def expected_tags(runs, a, b):
    """SURVEY.md App. B5 on a plain run list: first/last = 1-based indices of the runs containing a / b; decoding
    starts at run `first`, or ONE RUN LATE when first % 10 == 0; runs past the end read as value 0."""
    import bisect
    st = [r[0] for r in runs]
    first = bisect.bisect_right(st, a); last = bisect.bisect_right(st, b)
    shift = 1 if first % 10 == 0 else 0
    vals = [runs[i][1] if i < len(runs) else 0 for i in range(first - 1 + shift, last + shift)]
    return last - first + 1, sorted(set(vals)), first

def test_tag_query_equals_row_lookup(syn):
    """Both tag layouts against a brute-force lookup over the known per-row values, including the documented
    off-by-one of SURVEY.md §0.6."""
    s, _ = syn
    o = orc.Index(s["ri"], s["tags0"], 0); o1 = orc.Index(s["ri"], s["tags1"], 1)
    runs = tag_runs(s["row_val"]); assert len(runs) == s["n_tag_runs"]
    n = o.n; rng = np.random.default_rng(5); quirk = 0
    for _ in range(1500):
        a = int(rng.integers(0, n - 1)); b = min(n - 1, a + int(rng.integers(0, 80)))
        nr, want, first = expected_tags(runs, a, b)
        for ix in (o, o1):
            got_nr, tv = ix.tag_query(a, b)
            assert got_nr == nr and [int(x) for x in tv] == want, (a, b)
        if first % 10 == 0:
            quirk += 1
        else:
            assert want == sorted(set(int(v) for v in s["row_val"][a:b + 1]))    # the sane case: exactly the rows' values
    assert quirk > 50

def test_first_mod_10_quirk_is_reproduced(syn):
    """SURVEY.md §0.6 made explicit: a single-row query inside run number 10, 20, ... returns the NEXT run's value."""
    s, _ = syn
    o = orc.Index(s["ri"], s["tags0"], 0)
    runs = tag_runs(s["row_val"])
    for k in (10, 20, 50, 100):
        p = runs[k - 1][0]                                       # first row of the k-th run (1-based)
        nr, tv = o.tag_query(p, p)
        assert nr == 1 and [int(x) for x in tv] == [runs[k][1]] and runs[k][1] != runs[k - 1][1]
    p = runs[10][0]; nr, tv = o.tag_query(p, p)                  # run 11: sane
    assert [int(x) for x in tv] == [runs[10][1]]
