"""Synthetic pangenome workloads for bench.py / tests (BASELINE.json configs 3-5).

DATA PREP, not the product path: a random reference with a few N stretches, haplotypes carrying SNPs and short
indels at shared sites (allele frequency ~ Beta(0.2, 0.2)), their reverse complements, the multi-string BWT
(S1 $1 S2 $2 ... with $1 < $2 < ... < 'A', every $ written '\\n'), per-row graph positions (nodes of 32 bp on the
reference coordinate) and error-carrying reads. Torch is used here as plumbing to do this on the GPU quickly
(suffix sorting by prefix doubling); the files that come out are the reference's own formats (.rl_bwt -> encoded
.ri via prep/, sdsl-compact tags) and are loaded through the same C ABI as any user's index.
"""
import time
import numpy as np
import torch

NL, A, C, G, N, T = 10, 65, 67, 71, 78, 84

def _lut_comp(dev):
    lut = torch.arange(256, dtype=torch.uint8, device=dev)
    for a, b in ((A, T), (C, G), (G, C), (T, A)):
        lut[a] = b
    return lut

def make_haplotypes(ref_len, n_hap, seed, dev, snp=1e-3, indel=1e-4, n_stretch=10, stretch=100):
    """-> (text uint8 [n], refpos int32 [n], rev uint8 [n], seq_starts int64 [2*n_hap+1]); sequences are hap0, rc(hap0), hap1, ..."""
    g = torch.Generator(device=dev); g.manual_seed(seed)
    acgt = torch.tensor([A, C, G, T], dtype=torch.uint8, device=dev)
    refc = torch.randint(0, 4, (ref_len,), generator=g, device=dev)
    ref = acgt[refc]
    for _ in range(n_stretch if ref_len > 20 * stretch else 1):
        p = int(torch.randint(0, max(1, ref_len - stretch), (1,), generator=g, device=dev))
        ref[p:p + min(stretch, ref_len // 20 + 1)] = N
    beta = torch.distributions.Beta(torch.tensor(0.2), torch.tensor(0.2))
    torch.manual_seed(seed)
    snp_site = torch.nonzero(torch.rand(ref_len, generator=g, device=dev) < snp).flatten()
    snp_alt = acgt[(refc[snp_site] + 1 + torch.randint(0, 3, (len(snp_site),), generator=g, device=dev)) % 4]
    snp_f = beta.sample((len(snp_site),)).to(dev)
    ind_site = torch.nonzero(torch.rand(ref_len, generator=g, device=dev) < indel).flatten()
    ind_len = torch.randint(1, 11, (len(ind_site),), generator=g, device=dev)
    ind_del = torch.rand(len(ind_site), generator=g, device=dev) < 0.5
    ind_f = beta.sample((len(ind_site),)).to(dev)
    comp = _lut_comp(dev)
    pieces, rp, rv, starts = [], [], [], [0]
    pos = torch.arange(ref_len, device=dev, dtype=torch.int64)
    nl = torch.tensor([NL], dtype=torch.uint8, device=dev); z32 = torch.zeros(1, dtype=torch.int32, device=dev); z8 = torch.zeros(1, dtype=torch.uint8, device=dev)
    for h in range(n_hap):
        seq = ref.clone()
        take = torch.rand(len(snp_site), generator=g, device=dev) < snp_f
        seq[snp_site[take]] = snp_alt[take]
        has = torch.rand(len(ind_site), generator=g, device=dev) < ind_f
        d = has & ind_del; i = has & ~ind_del
        delta = torch.zeros(ref_len + 16, dtype=torch.int32, device=dev)
        delta.index_add_(0, ind_site[d] + 1, torch.ones(int(d.sum()), dtype=torch.int32, device=dev))
        delta.index_add_(0, torch.clamp(ind_site[d] + 1 + ind_len[d], max=ref_len + 15), -torch.ones(int(d.sum()), dtype=torch.int32, device=dev))
        keep = (torch.cumsum(delta[:ref_len], 0) <= 0)
        ins = torch.zeros(ref_len, dtype=torch.int64, device=dev); ins[ind_site[i]] = ind_len[i]
        out_len = keep.to(torch.int64) + ins
        off = torch.cumsum(out_len, 0) - out_len
        total = int(out_len.sum())
        hs = torch.empty(total, dtype=torch.uint8, device=dev); hp = torch.empty(total, dtype=torch.int32, device=dev)
        hs[off[keep]] = seq[keep]; hp[off[keep]] = pos[keep].to(torch.int32)
        if int(i.sum()):
            isite = ind_site[i]; il = ind_len[i]
            rep = torch.repeat_interleave(torch.arange(len(isite), device=dev), il)
            within = torch.arange(len(rep), device=dev) - torch.repeat_interleave(torch.cumsum(il, 0) - il, il)
            dst = off[isite][rep] + keep[isite][rep].to(torch.int64) + within
            hs[dst] = acgt[torch.randint(0, 4, (len(rep),), generator=g, device=dev)]; hp[dst] = isite[rep].to(torch.int32)
        for s_, p_, r_ in ((hs, hp, 0), (comp[hs.long()].flip(0), hp.flip(0), 1)):
            pieces += [s_, nl]; rp += [p_, z32]; rv += [torch.full((len(s_),), r_, dtype=torch.uint8, device=dev), z8]
            starts.append(starts[-1] + len(s_) + 1)
    return torch.cat(pieces), torch.cat(rp), torch.cat(rv), torch.tensor(starts, dtype=torch.int64, device=dev)

def suffix_array(text, dev, log=None):
    """Prefix doubling with distinct, ordered sentinels: rank('\\n' of sequence i) = i, letters by byte value above them."""
    n = len(text)
    is_nl = text == NL
    n_seq = int(is_nl.sum())
    rank = torch.where(is_nl, torch.cumsum(is_nl.to(torch.int64), 0) - 1, text.to(torch.int64) + n_seq)
    def shifted(r, h):
        out = torch.zeros_like(r); out[: n - h] = r[h:]; return out
    h = 1
    perm = None
    while True:
        mx = int(rank.max()) + 1
        key = rank * mx + shifted(rank, h)
        skey, perm = torch.sort(key)
        del key
        flag = torch.ones(n, dtype=torch.int64, device=dev); flag[1:] = (skey[1:] != skey[:-1]).to(torch.int64)
        del skey
        nr = torch.cumsum(flag, 0) - 1
        del flag
        distinct = int(nr[-1]) + 1
        rank = torch.empty_like(nr); rank[perm] = nr
        del nr
        if log: log("  prefix doubling h=%d distinct=%d/%d" % (h, distinct, n))
        if distinct == n: break
        h *= 2
        if h > 2 * n: raise RuntimeError("suffix sorting did not converge")
    return perm

def build(ref_len, n_hap, seed=3, device=0, log=None):
    """-> dict(text (cuda uint8), seq_starts, rl_sym, rl_len, tag_val, tag_len (numpy), n, n_seq)."""
    dev = torch.device(device) if isinstance(device, str) else torch.device("cuda", device)
    t0 = time.time()
    text, refpos, rev, starts = make_haplotypes(ref_len, n_hap, seed, dev)
    n = len(text)
    if log: log("synth: %d haplotypes x %d bp (+ reverse complements): n = %d (%.1fs)" % (n_hap, ref_len, n, time.time() - t0))
    sa = suffix_array(text, dev, log)
    prev = sa - 1; prev[prev < 0] = n - 1
    bwt = text[prev]; del prev
    sym, cnt = torch.unique_consecutive(bwt, return_counts=True); del bwt
    posval = ((refpos.to(torch.int64) // 32 + 1) << 11) | (rev.to(torch.int64) << 10) | (refpos.to(torch.int64) % 32)
    posval[text == NL] = 0                                     # endmarker rows carry 0 (merge_tags.cpp:619-622)
    rows = posval[sa]; del posval, sa
    tv, tl = torch.unique_consecutive(rows, return_counts=True); del rows
    if log: log("synth: r = %d BWT runs, %d tag runs (%.1fs)" % (len(sym), len(tv), time.time() - t0))
    out = dict(text=text, seq_starts=starts.cpu().numpy(), rl_sym=sym.cpu().numpy(), rl_len=cnt.cpu().numpy().astype(np.uint64),
               tag_val=tv.cpu().numpy().astype(np.uint64), tag_len=tl.cpu().numpy().astype(np.uint64), n=n, n_seq=2 * n_hap)
    if dev.type == "cuda": torch.cuda.empty_cache()
    return out

def sample_reads(text, seq_starts, n_reads, read_len, sub, seed, dev, indel=0.0):
    """Uniform over valid start positions; i.i.d. substitutions (and 1-bp indels for the long-read config).
    -> cuda uint8 [n_reads, read_len]."""
    g = torch.Generator(device=dev); g.manual_seed(seed)
    st = torch.as_tensor(seq_starts, device=dev)
    lens = st[1:] - st[:-1] - 1
    margin = read_len // 8 + 16 if indel > 0 else 0
    valid = torch.clamp(lens - read_len - margin + 1, min=0)
    cum = torch.cumsum(valid, 0)
    u = (torch.rand(n_reads, generator=g, device=dev, dtype=torch.float64) * float(cum[-1])).to(torch.int64)
    sidx = torch.searchsorted(cum, u, right=True)
    start = st[:-1][sidx] + (u - (cum[sidx] - valid[sidx]))
    acgt = torch.tensor([A, C, G, T], dtype=torch.uint8, device=dev)
    out = torch.empty((n_reads, read_len), dtype=torch.uint8, device=dev)
    step = max(1, (1 << 28) // read_len)
    ar = torch.arange(read_len, device=dev)
    for a in range(0, n_reads, step):
        b = min(n_reads, a + step)
        if indel > 0:
            ev = torch.rand((b - a, read_len), generator=g, device=dev)
            dele = (ev < indel / 2).to(torch.int64); inse = (ev >= indel / 2) & (ev < indel)
            srcoff = torch.cumsum(1 + dele - inse.to(torch.int64), 1) - 1
            srcoff = torch.clamp(srcoff, min=0, max=read_len + margin - 1)
            r = text[start[a:b, None] + srcoff]
            r = torch.where(inse, acgt[torch.randint(0, 4, r.shape, generator=g, device=dev)], r)
        else:
            r = text[start[a:b, None] + ar[None, :]]
        m = torch.rand(r.shape, generator=g, device=dev) < sub
        ridx = torch.randint(0, 4, r.shape, generator=g, device=dev)
        alt = torch.where(acgt[ridx] == r, acgt[(ridx + 1) % 4], acgt[ridx])     # a substitution always changes the base
        out[a:b] = torch.where(m, alt, r)
    return out
