"""pangraph.py (data prep for BASELINE configs 3-5 at nominal size): the BWT and the tag runs it derives from the structure
of the variation graph must equal, symbol for symbol, those of the suffix-array construction over the explicit text."""
import numpy as np, pytest
from conftest import load_package

def explicit(pg, g):
    """Text (hap0, rc(hap0), hap1, ...; '\\n' after each) and the encoded graph position of every text position."""
    seqs, vals = [], []
    gm = pg.mirror(g)
    for h in range(g.H):
        c, rp = pg.haplotype_codes(g, h)
        cm, rpm = pg.haplotype_codes(gm, h)
        assert np.array_equal(cm, pg.COMP_CODE[c[::-1]])                   # the mirrored graph spells the reverse complement
        fwd_rp = rp.astype(np.int64)
        seqs.append(pg.BYTE_OF[c].tobytes()); vals.append(((fwd_rp // 32 + 1) << 11) | (fwd_rp % 32))
        rrp = fwd_rp[::-1]
        seqs.append(pg.BYTE_OF[cm].tobytes()); vals.append(((rrp // 32 + 1) << 11) | (1 << 10) | (rrp % 32))
    return seqs, vals

def reference_runs(prep, seqs, vals):
    text = b"".join(s + b"\n" for s in seqs)
    rl = prep.bwt_from_text(text)
    ri, info, rows = prep.rindex_from_rlbwt(rl, encoded=True, with_samples=False, want_rows=True)
    ml = info["max_length"]; sid = (rows // np.uint64(ml)).astype(np.int64); off = (rows % np.uint64(ml)).astype(np.int64)
    lens = np.array([len(s) for s in seqs]); val = np.zeros(len(rows), dtype=np.uint64)
    for s in range(len(seqs)):
        sel = (sid == s) & (off < lens[s]); val[sel] = vals[s][off[sel]].astype(np.uint64)
    sb, fb = np.frombuffer(rl[:16], dtype=np.uint64); rec = np.frombuffer(rl[16:], dtype=np.uint8).reshape(-1, int(sb + fb))
    ln = np.zeros(len(rec), dtype=np.int64)
    for i in range(int(fb)): ln |= rec[:, 1 + i].astype(np.int64) << (8 * i)
    bwt = np.repeat(rec[:, 0], ln)
    return bwt, val

def expand(sym, ln): return np.repeat(np.asarray(sym), np.asarray(ln).astype(np.int64))

@pytest.mark.parametrize("ref_len,n_hap,seed,snp,indel", [(3000, 4, 1, 1 / 150, 1 / 300), (6000, 7, 2, 1 / 100, 1 / 200), (4000, 70, 3, 1 / 120, 1 / 250),
                                                          (12000, 16, 4, 1e-3, 1e-4), (5000, 3, 5, 0.0, 1 / 100), (5000, 130, 6, 1 / 60, 1 / 400)])
def test_structural_bwt_equals_suffix_array_bwt(ref_len, n_hap, seed, snp, indel):
    pkg = load_package(); pg = pkg.pangraph; prep = pkg.prep
    g = pg.make_graph(ref_len, n_hap, seed, snp=snp, indel=indel, n_stretch=2)
    assert g.S > 5
    d = pg.build(ref_len, n_hap, graph=g, device="cpu")
    seqs, vals = explicit(pg, g)
    bwt, val = reference_runs(prep, seqs, vals)
    assert d["n"] == len(bwt) and d["n_seq"] == 2 * n_hap
    got = expand(d["rl_sym"], d["rl_len"])
    assert len(got) == len(bwt)
    bad = np.nonzero(got != bwt)[0]
    assert len(bad) == 0, "first BWT mismatch at row %d" % bad[0]
    gv = expand(d["tag_val"], d["tag_len"])
    bad = np.nonzero(gv != val)[0]
    assert len(bad) == 0, "first tag mismatch at row %d" % bad[0]
    assert (np.asarray(d["rl_sym"])[1:] != np.asarray(d["rl_sym"])[:-1]).all()          # maximal runs

def test_sampled_reads_come_from_the_haplotypes():
    """Error-free reads must occur in the explicit text (both strands, through SNPs / insertions / deletions); reads with
    substitutions differ from their origin in about the requested fraction of positions."""
    pkg = load_package(); pg = pkg.pangraph
    g = pg.make_graph(30000, 6, 9, snp=1 / 100, indel=1 / 300, n_stretch=2)
    _, strands = pg.prepare(30000, 6, graph=g, device="cpu")
    seqs, _ = explicit(pg, g); text = b"\n".join(seqs)
    r = pg.sample_reads(strands, 1500, 150, 0.0, seed=1).numpy()
    assert all(r[i].tobytes() in text for i in range(len(r)))
    assert len({r[i].tobytes() for i in range(len(r))}) > 1400
    e = pg.sample_reads(strands, 1500, 150, 0.02, seed=1).numpy()              # same seed: same origins, 2 % substitutions on top
    frac = (e != r).mean(); assert 0.012 < frac < 0.03
    long = pg.sample_reads(strands, 50, 2000, 0.03, seed=2, indel=0.02)
    assert long.shape == (50, 2000) and set(np.unique(long.numpy())) <= set(b"ACGTN")
