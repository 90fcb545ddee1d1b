"""Test datasets: the reference's shipped fixtures (tests/golden/) and small synthetic pangenomes built with
the prep tools (sigma = 6 bidirectional texts with aligned tags), cached per process."""
import functools, os
import numpy as np
from conftest import load_package

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

def gold(name):
    with open(os.path.join(GOLD, name), "rb") as f:
        return f.read()

_COMP = bytes.maketrans(b"ACGT", b"TGCA")
def revcomp(s: bytes) -> bytes:
    return s.translate(_COMP)[::-1]

def synth_haplotypes(ref_len, n_hap, seed, snp=1 / 200, indel=1 / 1000, n_N=2):
    """Random reference with a few N stretches; haplotypes carry SNPs / short indels at shared sites.
    Returns (list of haplotype bytes, list of per-haplotype arrays mapping haplotype offset -> reference offset)."""
    rng = np.random.default_rng(seed)
    ref = rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), ref_len)
    for _ in range(n_N):
        p = int(rng.integers(0, max(1, ref_len - 12)))
        ref[p:p + int(rng.integers(1, 10))] = ord("N")
    sites = np.nonzero(rng.random(ref_len) < snp)[0]
    alt = rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), len(sites))
    freq = rng.beta(0.4, 0.4, len(sites))
    isites = np.nonzero(rng.random(ref_len) < indel)[0]
    ilen = rng.integers(1, 6, len(isites)); idel = rng.random(len(isites)) < 0.5; ifreq = rng.beta(0.4, 0.4, len(isites))
    haps, maps = [], []
    for h in range(n_hap):
        seq = ref.copy()
        take = rng.random(len(sites)) < freq
        seq[sites[take]] = alt[take]
        keep = np.ones(ref_len, dtype=bool)
        ins_after = {}
        for s, l, d, f in zip(isites, ilen, idel, ifreq):
            if rng.random() < f:
                if d: keep[s:s + l] = False
                else: ins_after[int(s)] = rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), int(l))
        out, omap = [], []
        last = 0
        for s in sorted(ins_after):
            seg = np.arange(last, s + 1)[keep[last:s + 1]]
            out.append(seq[seg]); omap.append(seg)
            out.append(ins_after[s]); omap.append(np.full(len(ins_after[s]), s))
            last = s + 1
        seg = np.arange(last, ref_len)[keep[last:]]
        out.append(seq[seg]); omap.append(seg)
        haps.append(np.concatenate(out).tobytes()); maps.append(np.concatenate(omap))
    return haps, maps

@functools.lru_cache(maxsize=None)
def synth_index(ref_len=3000, n_hap=4, seed=7, encoded=True, n_N=2):
    """-> dict(text, ri, tags0 (sdsl-compact), tags1 (ByteCode+len9), info)."""
    prep = load_package().prep
    haps, maps = synth_haplotypes(ref_len, n_hap, seed, n_N=n_N)
    seqs, smaps, srev = [], [], []
    for h, m in zip(haps, maps):
        seqs.append(h); smaps.append(m); srev.append(0)
        seqs.append(revcomp(h)); smaps.append(m[::-1]); srev.append(1)
    text = b"".join(s + b"\n" for s in seqs)
    rl = prep.bwt_from_text(text)
    ri, info, rows = prep.rindex_from_rlbwt(rl, encoded=encoded, with_samples=False, want_rows=True)
    ml = info["max_length"]
    seq_id = (rows // np.uint64(ml)).astype(np.int64); off = (rows % np.uint64(ml)).astype(np.int64)
    lens = np.array([len(s) for s in seqs])
    val = np.zeros(len(rows), dtype=np.uint64)
    for s in range(len(seqs)):
        sel = (seq_id == s) & (off < lens[s])
        refpos = smaps[s][off[sel]]
        val[sel] = ((refpos // 32 + 1).astype(np.uint64) << np.uint64(11)) | np.uint64(srev[s] << 10) | (refpos % 32).astype(np.uint64)
    val[(off >= lens[seq_id])] = 0            # endmarker rows (merge_tags.cpp:619-622)
    tags0, nr0 = prep.tags_from_rows(val, 0)
    tags1, nr1 = prep.tags_from_rows(val, 1)
    return dict(text=text, seqs=seqs, ri=ri, tags0=tags0, tags1=tags1, info=info, rl_bwt=rl, n_tag_runs=nr0, row_val=val)

def sample_reads(text, n_reads, read_len, sub, seed, indel=0.0):
    prep = load_package().prep
    r = prep.sample_reads(text, n_reads, read_len, sub=sub, indel=indel, seed=seed)
    off = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(read_len)
    return r.reshape(-1).copy(), off

def ragged_reads(text, seed, n=200):
    """Edge cases: empty reads, reads shorter than min_len, reads with N / lower case / NUL / high bytes, ragged lengths."""
    rng = np.random.default_rng(seed)
    lines = [l for l in text.split(b"\n") if len(l) > 300]
    out = [b"", b"A", b"ACGT", b"N" * 30, b"acgtacgtacgtacgtacgtacgt", bytes([0, 65, 67, 200, 71, 84] * 6)]
    for _ in range(n):
        L = lines[int(rng.integers(len(lines)))]; ln = int(rng.integers(1, 300)); st = int(rng.integers(0, len(L) - ln))
        r = bytearray(L[st:st + ln])
        for i in range(len(r)):
            if rng.random() < 0.02: r[i] = b"ACGTN"[int(rng.integers(5))]
        out.append(bytes(r))
    return out
