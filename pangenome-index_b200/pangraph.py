"""Exact BWT + tag runs of a synthetic pangenome WITHOUT suffix-sorting its text (BASELINE.json configs 3-5 at nominal size).

DATA PREP, not the product path. synth.py sorts all n suffixes (prefix doubling, ~50 n bytes): it stops near n = 3e9. The
collections of configs 3-5 (64 x 50 Mbp: n = 6.4e9; 256 x 150 Mbp: n = 7.7e10) are haplotypes of ONE variation graph, and
that structure gives the BWT directly, in O(reference length + sites x haplotypes) work:

  * a COLUMN = one position of the graph (a reference base, the alternative base of a SNP, one inserted base); all haplotypes
    that pass a column share its tag (graph position), so suffixes are handled per column, never per text position;
  * the suffixes of one column are ordered by the haplotypes' FUTURE allele choices. With sites kept >= GAP bases apart and the
    two alleles of a site decided against each other within GAP characters (checked at generation), that order is the
    positional BWT order O_j of the next site j: O_j = stable partition of O_{j+1} by the allele at j (smaller continuation
    first). A column's haplotypes split into CLASSES by their alleles at the few sites within reach of the next W characters;
    a class is a contiguous block of O_j, and all its suffixes share their first W characters;
  * classes of DIFFERENT columns are ordered by those W = 42 characters (126-bit keys); the reference is random, so two
    columns never share 42 characters -- this is CHECKED after the sort (a tie between columns raises), which makes the
    result exact rather than probable;
  * the BWT symbol of a suffix is the character before its column: the same for a whole class unless a site ends right
    before the column, in which case the class is expanded into its haplotypes in O order.
The reverse-complement strands are the same construction on the mirrored graph. Sorting ~1.2 records per graph column
replaces sorting n suffixes (config 4: 3.5e8 records instead of 7.7e10 suffixes).

tests/test_pangraph.py checks the result symbol for symbol against the suffix-array construction on explicit texts.
Everything is torch tensor plumbing on `dev` (CPU in the tests, the B200 in bench.py); no kernel of the product is involved.
"""
import numpy as np
import torch

W = 42            # characters of a sort key: two int64 of 21 x 3 bits
HALF = 21
GAP = 16          # minimum distance between two sites (and maximum decision depth of a site's two alleles)
KWIN = 6          # sites that can lie within reach of one window
REACH = W + 12 * KWIN   # reference positions W characters can span: every carried deletion (<= 12 bases) extends them
END_MARGIN = 320  # no sites this close to either end of the reference
NL, A, C, G, N, T = 10, 65, 67, 71, 78, 84
CODE_OF = {NL: 0, A: 1, C: 2, G: 3, N: 4, T: 5}      # nuc order = byte order (include/pangenome_index/utils.hpp:11)
BYTE_OF = np.array([NL, A, C, G, N, T, 0, 0], dtype=np.uint8)
COMP_CODE = np.array([0, 5, 3, 2, 4, 1, 0, 0], dtype=np.uint8)    # complement on codes: A<->T, C<->G, N, $ fixed
SNP, DEL, INS = 0, 1, 2

# --------------------------------------------------------------------------------------------------------------- the graph
class Graph:
    """Reference + biallelic sites + haplotype allele matrix, in FORWARD coordinates (numpy, host)."""
    def __init__(self, ref_codes, kind, pos, length, alt_codes, ins_off, ins_codes, alleles):
        self.ref = ref_codes            # uint8 [L] codes 1..5
        self.kind, self.pos, self.len = kind, pos, length          # [S]: SNP / DEL / INS, anchor position, indel length (SNP: 1)
        self.alt = alt_codes            # [S] code of the SNP's alternative base (0 for indels)
        self.ins_off, self.ins = ins_off, ins_codes                 # [S + 1] offsets into the inserted bases
        self.alleles = alleles          # uint8 [H, S]: 1 = the haplotype carries the alternative allele
        self.L = len(ref_codes); self.S = len(pos); self.H = alleles.shape[0]

def _varstart_end(kind, pos, length):
    vs = np.where(kind == SNP, pos, pos + 1)
    ve = np.where(kind == SNP, pos + 1, np.where(kind == DEL, pos + length + 1, pos + 1))
    return vs.astype(np.int64), ve.astype(np.int64)

def _lcp_np(a, b, cap):
    n = min(len(a), len(b), cap); d = 0
    while d < n and a[d] == b[d]: d += 1
    return d

def make_graph(ref_len, n_hap, seed, snp=1e-3, indel=1e-4, n_stretch=10, stretch=24, beta=0.2):
    """Random ACGT reference with a few N stretches (sigma = 6), SNPs at rate `snp` and indels of 1..10 bases at rate `indel`
    per base, allele frequencies ~ Beta(beta, beta); sites closer than GAP to their predecessor, inside / next to an N
    stretch, near the ends, or whose two alleles are not decided within GAP characters are dropped."""
    rng = np.random.default_rng(seed)
    ref = rng.integers(1, 5, ref_len).astype(np.uint8); ref[ref == 4] = 5          # codes 1, 2, 3, 5 = A, C, G, T
    isN = np.zeros(ref_len + 1, dtype=bool)
    if ref_len > 40 * stretch:
        for _ in range(n_stretch):
            p = int(rng.integers(END_MARGIN, ref_len - END_MARGIN - stretch)); ref[p:p + stretch] = 4; isN[max(0, p - 12):p + stretch + 12] = True
    n_snp = rng.binomial(ref_len, snp); n_ind = rng.binomial(ref_len, indel)
    pos = np.concatenate([rng.integers(0, ref_len, n_snp), rng.integers(0, ref_len, n_ind)])
    kind = np.concatenate([np.full(n_snp, SNP), np.where(rng.random(n_ind) < 0.5, DEL, INS)]).astype(np.int64)
    length = np.concatenate([np.ones(n_snp, dtype=np.int64), rng.integers(1, 11, n_ind)])
    o = np.argsort(pos, kind="stable"); pos, kind, length = pos[o], kind[o], length[o]
    vs, ve = _varstart_end(kind, pos, length)
    keep = (vs >= END_MARGIN) & (ve <= ref_len - END_MARGIN) & ~isN[np.minimum(pos, ref_len)] & ~isN[np.minimum(ve, ref_len)]
    pos, kind, length, vs, ve = pos[keep], kind[keep], length[keep], vs[keep], ve[keep]
    sel = []; last_end = -10 ** 9                         # greedy thinning: GAP bases between one site's end and the next one's start
    for i in range(len(pos)):
        if vs[i] - last_end < GAP + 2 or (sel and vs[i] - pos[sel[-1]] < GAP): continue      # (+ 2: a deletion may be lengthened below)
        if len(sel) >= KWIN and vs[i] - vs[sel[-KWIN]] <= REACH + 4: continue                     # at most KWIN sites within reach of one window
        sel.append(i); last_end = ve[i]
    sel = np.array(sel, dtype=np.int64); pos, kind, length = pos[sel], kind[sel], length[sel]
    S = len(pos)
    acgt = np.array([1, 2, 3, 5], dtype=np.uint8); idx_of = np.array([0, 0, 1, 2, 0, 3], dtype=np.int64)
    alt = np.where(kind == SNP, acgt[(idx_of[ref[pos]] + 1 + rng.integers(0, 3, S)) % 4], 0).astype(np.uint8)
    ins_len = np.where(kind == INS, length, 0); ins_off = np.concatenate([[0], np.cumsum(ins_len)]).astype(np.int64)
    ins = acgt[rng.integers(0, 4, int(ins_off[-1]))]
    # Indels are kept NORMALISED at both ends: the base an allele removes / adds next to the anchor differs from the base that
    # would take its place (deletion: ref[pos + len] != ref[pos] and ref[pos + 1] != ref[pos + len + 1]; insertion X: X[-1] !=
    # ref[pos] and X[0] != ref[pos + 1]). Otherwise the carriers' suffixes from one column and the non-carriers' suffixes
    # from a shifted column are the same string for thousands of characters (the usual left / right shift ambiguity of an
    # indel) and two different columns would interleave haplotype by haplotype.
    ok = np.ones(S, dtype=bool)
    for i in np.nonzero(kind == DEL)[0]:
        p = int(pos[i]); ln = int(length[i])
        while ln <= 12 and (ref[p + ln] == ref[p] or ref[p + 1] == ref[p + ln + 1]): ln += 1
        if ln > 12: ok[i] = False
        length[i] = min(ln, 12)
    for i in np.nonzero(kind == INS)[0]:
        p = int(pos[i]); a, b = int(ins_off[i]), int(ins_off[i + 1])
        for _ in range(16):
            if ins[b - 1] != ref[p] and ins[a] != ref[p + 1]: break
            ins[a:b] = acgt[rng.integers(0, 4, b - a)]
        else: ok[i] = False
    vs, ve = _varstart_end(kind, pos, length)
    ok[1:] &= (vs[1:] - ve[:-1] >= GAP)                   # a lengthened deletion may have come too close to its successor
    keep_ins = np.concatenate([np.arange(ins_off[i], ins_off[i + 1]) for i in np.nonzero(ok)[0]] + [np.zeros(0, dtype=np.int64)]).astype(np.int64)
    pos, kind, length, alt, ins = pos[ok], kind[ok], length[ok], alt[ok], ins[keep_ins]
    ins_len = np.where(kind == INS, length, 0); ins_off = np.concatenate([[0], np.cumsum(ins_len)]).astype(np.int64)
    S = len(pos)
    g = Graph(ref, kind, pos.astype(np.int64), length.astype(np.int64), alt, ins_off, ins, np.zeros((1, S), dtype=np.uint8))
    # the two alleles of an indel must be decided against each other well within GAP characters, on both strands
    ok = (_decision_depth(g) < GAP - 3) & (_decision_depth(mirror(g))[::-1] < GAP - 3)
    if not ok.all():
        keep_ins = np.concatenate([np.arange(ins_off[i], ins_off[i + 1]) for i in np.nonzero(ok)[0]] + [np.zeros(0, dtype=np.int64)]).astype(np.int64)
        pos, kind, length, alt, ins = pos[ok], kind[ok], length[ok], alt[ok], ins[keep_ins]
        ins_len = np.where(kind == INS, length, 0); ins_off = np.concatenate([[0], np.cumsum(ins_len)]).astype(np.int64)
        S = len(pos)
    freq = rng.beta(beta, beta, S)
    alleles = (rng.random((n_hap, S)) < freq[None, :]).astype(np.uint8)
    return Graph(ref, kind, pos.astype(np.int64), length.astype(np.int64), alt, ins_off, ins, alleles)

def _continuations(g, j, depth):
    """(alternative, reference) allele continuations of an indel from its variable start, `depth` characters of reference behind."""
    p, k, ln = int(g.pos[j]), int(g.kind[j]), int(g.len[j])
    if k == DEL: return g.ref[p + ln + 1:p + ln + 1 + depth], g.ref[p + 1:p + 1 + depth]
    return np.concatenate([g.ins[g.ins_off[j]:g.ins_off[j + 1]], g.ref[p + 1:p + 1 + depth]]), g.ref[p + 1:p + 1 + depth]

def _decision_depth(g):
    """Characters the two alleles of every site share from its variable start (SNP: 0)."""
    d = np.zeros(g.S, dtype=np.int64)
    for j in np.nonzero(g.kind != SNP)[0]:
        a, b = _continuations(g, j, 2 * GAP); d[j] = _lcp_np(a, b, 2 * GAP)
    return d

def mirror(g):
    """The graph read on the reverse-complement strand: haplotype h of mirror(g) spells the reverse complement of haplotype h."""
    L = g.L
    ref = COMP_CODE[g.ref[::-1]].copy()
    pos = np.where(g.kind == SNP, L - 1 - g.pos, np.where(g.kind == DEL, L - 2 - g.pos - g.len, L - 2 - g.pos))[::-1].copy()
    kind = g.kind[::-1].copy(); length = g.len[::-1].copy(); alt = COMP_CODE[g.alt[::-1]].copy()
    ins_len = np.where(kind == INS, length, 0); ins_off = np.concatenate([[0], np.cumsum(ins_len)]).astype(np.int64)
    pieces = [COMP_CODE[g.ins[g.ins_off[j]:g.ins_off[j + 1]][::-1]] for j in range(g.S - 1, -1, -1)]
    ins = np.concatenate(pieces) if pieces else g.ins[:0]
    m = Graph(ref, kind, pos.astype(np.int64), length, alt, ins_off, ins.astype(np.uint8), g.alleles[:, ::-1].copy())
    m.orig_anchor = g.pos[::-1].copy()          # forward-strand anchor of every site (tags of inserted bases)
    return m

def haplotype_codes(g, h):
    """Explicit sequence of haplotype h (codes) and the forward... graph-local reference position of every base (tests only)."""
    out, rp = [], []; q = 0
    for j in range(g.S):
        p, k, ln, a = int(g.pos[j]), int(g.kind[j]), int(g.len[j]), int(g.alleles[h, j])
        if k == SNP:
            out.append(g.ref[q:p]); rp.append(np.arange(q, p)); out.append(np.array([g.alt[j] if a else g.ref[p]], dtype=np.uint8)); rp.append(np.array([p])); q = p + 1
        elif k == DEL:
            out.append(g.ref[q:p + 1]); rp.append(np.arange(q, p + 1)); q = p + 1 + (ln if a else 0)
        else:
            out.append(g.ref[q:p + 1]); rp.append(np.arange(q, p + 1)); q = p + 1
            if a: out.append(g.ins[g.ins_off[j]:g.ins_off[j + 1]]); rp.append(np.full(ln, p))
    out.append(g.ref[q:]); rp.append(np.arange(q, g.L))
    return np.concatenate(out), np.concatenate(rp)

# --------------------------------------------------------------------------------------------------------------- PBWT orders
def _allele_order(g):
    """ord[j] = 1 when the ALTERNATIVE allele's continuation is lexicographically smaller than the reference allele's."""
    o = (g.alt < g.ref[np.minimum(g.pos, g.L - 1)]).astype(np.uint8)
    for j in np.nonzero(g.kind != SNP)[0]:
        a, b = _continuations(g, j, 2 * GAP); d = _lcp_np(a, b, 2 * GAP)
        if d >= min(len(a), len(b)): raise RuntimeError("pangraph: alleles of site %d are not decided (generator invariant broken)" % j)
        o[j] = a[d] < b[d]
    return o

def pbwt_orders(g):
    """order[j] = haplotypes sorted by their suffix from the variable start of site j (j = S: by sequence number).
    Returns the permuted allele tables PA[j, t + 1, word] (t = -1 .. KWIN - 1): bit i of PA[j, t + 1] = allele at site j + t of the
    haplotype of rank i in order[j], as ceil(H / 64) int64 words; plus order itself."""
    H, S = g.H, g.S; ordj = _allele_order(g)
    order = np.zeros((S + 1, H), dtype=np.int32); cur = np.arange(H, dtype=np.int32); order[S] = cur
    al = g.alleles
    for j in range(S - 1, -1, -1):
        b = al[cur, j]
        first = b == ordj[j]                      # the smaller allele first, stable
        cur = np.concatenate([cur[first], cur[~first]]); order[j] = cur
    HW = (H + 63) // 64
    PA = np.zeros((S + 1, KWIN + 1, HW), dtype=np.int64)
    alp = np.concatenate([np.zeros((H, 1), dtype=np.uint8), al, np.zeros((H, KWIN + 1), dtype=np.uint8)], axis=1)   # column s + 1 = site s
    weights = (np.uint64(1) << np.arange(64, dtype=np.uint64))
    for t in range(-1, KWIN):
        cols = np.arange(S + 1) + t + 1                                      # site j + t -> padded column
        bits = alp[order, cols[:, None]]                                     # [S + 1, H]
        pad = np.zeros((S + 1, HW * 64), dtype=np.uint64); pad[:, :H] = bits
        PA[:, t + 1, :] = (pad.reshape(S + 1, HW, 64) * weights[None, None, :]).sum(axis=2, dtype=np.uint64).view(np.int64)
    return order, PA

# --------------------------------------------------------------------------------------------------------------- bit helpers (int64 tensors)
def _popcount64(x):
    m1, m2, m4 = 0x5555555555555555, 0x3333333333333333, 0x0F0F0F0F0F0F0F0F
    x = x - ((x >> 1) & m1)                       # arithmetic shift: the mask clears the smeared sign bits
    x = (x & m2) + ((x >> 2) & m2)
    x = (x + ((x >> 4) & 0x0FFFFFFFFFFFFFFF)) & m4
    return ((x * 0x0101010101010101) >> 56) & 0xFF
def _ctz64(x):
    """Index of the lowest set bit (64 for 0)."""
    low = x & (-x)
    return torch.where(x == 0, torch.full_like(x, 64), _popcount64(low - 1))

# --------------------------------------------------------------------------------------------------------------- one strand
class Strand:
    """Device-side arrays of one graph (forward or mirrored)."""
    def __init__(self, g, model, dev, orig_L, orig_anchor=None):
        t = lambda a, dt=torch.int64: torch.as_tensor(np.ascontiguousarray(a), device=dev).to(dt)
        self.g, self.model, self.dev, self.L, self.S, self.H = g, model, dev, g.L, g.S, g.H
        self.HW = (g.H + 63) // 64
        self.ref = torch.cat([t(g.ref, torch.uint8), torch.zeros(W + 16, dtype=torch.uint8, device=dev)])      # code 0 ('$') behind the end
        vs, ve = _varstart_end(g.kind, g.pos, g.len)
        big = np.int64(1) << 40
        self.vs = t(np.concatenate([vs, [big] * (KWIN + 2)])); self.ve = t(np.concatenate([ve, [big] * (KWIN + 2)]))
        self.kind = t(np.concatenate([g.kind, [SNP] * (KWIN + 2)])); self.pos = t(np.concatenate([g.pos, [big] * (KWIN + 2)]))
        self.len = t(np.concatenate([g.len, [0] * (KWIN + 2)])); self.alt = t(np.concatenate([g.alt, [0] * (KWIN + 2)]), torch.uint8)
        self.ins_off = t(np.concatenate([g.ins_off, [g.ins_off[-1]] * (KWIN + 2)])); self.ins = torch.cat([t(g.ins, torch.uint8), torch.zeros(16, dtype=torch.uint8, device=dev)])
        order, PA = pbwt_orders(g)
        self.order = order; self.PA = t(PA)                                # [S + 1, KWIN + 1, HW]
        full = np.zeros(self.HW, dtype=np.uint64)
        for w in range(self.HW):
            nb = min(64, g.H - 64 * w); full[w] = np.uint64(0xFFFFFFFFFFFFFFFF) if nb == 64 else (np.uint64(1) << np.uint64(nb)) - np.uint64(1)
        self.full = t(full.view(np.int64))
        self.orig_L = orig_L
        # forward-strand reference position of a column (tags): backbone c -> c or L - 1 - c; inserted bases -> the site's forward anchor
        self.anchor_fwd = t(np.concatenate([g.pos if model == 0 else orig_anchor, [0] * (KWIN + 2)]))

    def tag_of_refpos(self, rp):
        return ((rp // 32 + 1) << 11) | (self.model << 10) | (rp % 32)
    def fwd_pos(self, c):
        return c if self.model == 0 else (self.orig_L - 1 - c)

    # ---- W characters along a path. q: next reference position; jn: first site with varstart >= q that is still ahead;
    # ins_site / ins_off: inside an insertion; bit_of(t, site) -> allele (0/1 tensor) of the path at `site` (t = site - j0).
    def walk(self, q, jn, ins_site, ins_off, bit_of, steps, tail=None):
        n = len(q); dev = self.dev
        out = torch.zeros((n, steps), dtype=torch.uint8, device=dev)
        q = q.clone(); jn = jn.clone(); ins_site = ins_site.clone(); ins_off = ins_off.clone()
        ended = torch.zeros(n, dtype=torch.int64, device=dev)        # 0: running; k > 0: k - 1 tail characters emitted
        for i in range(steps):
            in_ins = ins_site >= 0
            trig = (~in_ins) & (self.vs[jn] == q)
            kd = self.kind[jn]
            b = torch.where(trig, bit_of(jn), torch.zeros_like(q)) if bool(trig.any()) else torch.zeros_like(q)
            carried = trig & (b == 1)
            # deletion carried: skip the deleted bases first, then read the reference
            q = torch.where(carried & (kd == DEL), q + self.len[jn], q)
            enter = carried & (kd == INS)
            ins_site = torch.where(enter, jn, ins_site); ins_off = torch.where(enter, torch.zeros_like(ins_off), ins_off)
            now_ins = ins_site >= 0
            ch_ins = self.ins[(self.ins_off[torch.clamp(ins_site, min=0)] + ins_off).clamp(max=len(self.ins) - 1)]
            ch_ref = self.ref[q.clamp(max=self.L + W)]
            ch = torch.where(now_ins, ch_ins, torch.where(carried & (kd == SNP), self.alt[jn], ch_ref))
            # end of the sequence: '$', then the sequence number in five 3-bit characters, then zeros
            at_end = (~now_ins) & (q >= self.L)
            if tail is not None:
                k = ended
                tch = torch.where(k == 0, torch.zeros_like(k), torch.where(k <= 5, (tail >> (3 * (5 - k)).clamp(min=0)) & 7, torch.zeros_like(k)))
                ch = torch.where(at_end, tch.to(torch.uint8), ch)
                ended = torch.where(at_end, ended + 1, ended)
            else:
                ch = torch.where(at_end, torch.zeros_like(ch), ch)
            out[:, i] = ch
            # advance
            ins_off = torch.where(now_ins, ins_off + 1, ins_off)
            leave = now_ins & (ins_off >= self.len[torch.clamp(ins_site, min=0)])
            ins_site = torch.where(leave, torch.full_like(ins_site, -1), ins_site)
            q = torch.where(now_ins, q, q + 1)
            jn = torch.where(trig, jn + 1, jn)
        return out

def _pack_keys(codes):
    """[n, 42] codes -> (hi, lo) int64 of 21 x 3 bits each."""
    sh = (3 * (HALF - 1 - torch.arange(HALF, device=codes.device))).to(torch.int64)
    c = codes.to(torch.int64)
    return (c[:, :HALF] << sh).sum(dim=1), (c[:, HALF:] << sh).sum(dim=1)

def _ref_keys(st):
    """(hi, lo) of the pure reference path from every backbone column, by doubling (no per-column loop)."""
    k1 = st.ref.to(torch.int64)
    n = st.L + 1
    def at(a, o): return a[o:o + n + W]
    pad = torch.cat([k1, torch.zeros(3 * W, dtype=torch.int64, device=st.dev)])
    k2 = (pad[:-1] << 3) | pad[1:]
    k4 = (k2[:-2] << 6) | k2[2:]
    k5 = (k4[:len(pad) - 4] << 3) | pad[4:]
    k10 = (k5[:-5] << 15) | k5[5:]
    k20 = (k10[:-10] << 30) | k10[10:]
    k21 = (k20[:len(pad) - 20] << 3) | pad[20:len(pad)][:len(k20)]
    hi = k21[:st.L]; lo = k21[HALF:HALF + st.L]
    return hi, lo

def strand_records(st, log=None):
    """All (column, class) records of one strand: dict of equal-length tensors
         hi, lo  sort key (first W characters)      sub  column id << 10 | rank inside the column
         ch      BWT symbol code                    cnt  number of suffixes (rows)     tag  encoded graph position"""
    dev, L, S, H, HW = st.dev, st.L, st.S, st.H, st.HW
    i64 = torch.int64
    rec = {k: [] for k in ("hi", "lo", "sub", "ch", "cnt", "tag")}
    def emit(hi, lo, col, rank, ch, cnt, tag):
        rec["hi"].append(hi); rec["lo"].append(lo); rec["sub"].append(((col + (st.model << 40)) << 10) | rank); rec["ch"].append(ch.to(torch.uint8)); rec["cnt"].append(cnt.to(torch.int32)); rec["tag"].append(tag)
    ref_hi, ref_lo = _ref_keys(st)
    END0 = L - (W + 64)                                  # columns from here on see the end of the sequence: one class per haplotype

    def masks_for(jn, forced_site_bit, pattern, k):
        """Class mask (HW words, in order[jn] bit order): haplotypes passing the column (forced allele at site jn - 1, or -1 = all)
        whose alleles at sites jn .. jn + k - 1 spell `pattern` (bit t = allele at jn + t)."""
        m = st.full[None, :].expand(len(jn), HW).clone()
        pa = st.PA[jn]                                   # [n, KWIN + 1, HW]
        prev = pa[:, 0, :]
        m = torch.where((forced_site_bit == 0)[:, None], m & ~prev, m)
        m = torch.where((forced_site_bit == 1)[:, None], m & prev, m)
        for t in range(KWIN):
            use = (t < k)[:, None]; v = pa[:, t + 1, :]; bit = ((pattern >> t) & 1)[:, None]
            m = torch.where(use, torch.where(bit == 1, m & v, m & ~v & st.full[None, :]), m)
        return m & st.full[None, :]
    def count_rank(m):
        cnt = torch.zeros(len(m), dtype=i64, device=dev); rank = torch.full((len(m),), 64 * HW, dtype=i64, device=dev)
        for w in range(HW - 1, -1, -1):
            cnt += _popcount64(m[:, w]); z = _ctz64(m[:, w]); rank = torch.where(m[:, w] != 0, z + 64 * w, rank)
        return cnt, rank

    def classes(col_id, q0, jn, forced, first_code, ins_site0, ins_off0, prev_ch, dep_site, tag):
        """Enumerate the classes of a set of columns and emit their records.
        q0: reference position where reading continues (for a backbone column: the column itself); jn: first site ahead of it;
        forced: allele forced at site jn - 1 (-1 none); first_code: the column's own character when it is not read from q0
        (SNP alternative base; 0 = read everything by walking); dep_site: >= 0 when the BWT symbol depends on the allele at
        site jn - 1 (the column right behind a site): the class is expanded into its haplotypes."""
        n = len(col_id)
        if n == 0: return
        k = torch.zeros(n, dtype=i64, device=dev)
        for t in range(KWIN):
            k += ((st.vs[jn + t] <= q0 + REACH) & (k == t)).to(i64)
        if bool((st.vs[jn + KWIN] <= q0 + REACH).any()): raise RuntimeError("pangraph: more than %d sites within reach of one window" % KWIN)
        # the common case -- no site within reach, every haplotype passes, one BWT symbol: one record straight from the reference keys
        simple = (k == 0) & (forced < 0) & (dep_site < 0) & (first_code == 0) & (ins_site0 < 0)
        if bool(simple.any()):
            qs = q0[simple]
            emit(ref_hi[qs], ref_lo[qs], col_id[simple], torch.zeros_like(qs), prev_ch[simple], torch.full_like(qs, H), tag[simple])
            rest = ~simple
            col_id, q0, jn, forced, first_code, ins_site0, ins_off0, prev_ch, dep_site, tag, k = (x[rest] for x in (col_id, q0, jn, forced, first_code, ins_site0, ins_off0, prev_ch, dep_site, tag, k))
            n = len(col_id)
            if n == 0: return
        npat = (1 << k)
        base = torch.cumsum(npat, 0) - npat; tot = int(npat.sum())
        owner = torch.repeat_interleave(torch.arange(n, device=dev), npat)
        pattern = torch.arange(tot, device=dev) - base[owner]
        jn_p, k_p, forced_p = jn[owner], k[owner], forced[owner]
        m = masks_for(jn_p, forced_p, pattern, k_p)
        cnt, rank = count_rank(m)
        keep = cnt > 0
        owner, pattern, jn_p, k_p, m, cnt, rank = owner[keep], pattern[keep], jn_p[keep], k_p[keep], m[keep], cnt[keep], rank[keep]
        # keys: the pure reference path where nothing is carried, a walk otherwise
        plain = (pattern == 0) & (first_code[owner] == 0) & (ins_site0[owner] < 0)
        hi = torch.zeros(len(owner), dtype=i64, device=dev); lo = torch.zeros_like(hi)
        qo = q0[owner]
        hi[plain] = ref_hi[qo[plain].clamp(max=L - 1)]; lo[plain] = ref_lo[qo[plain].clamp(max=L - 1)]
        wsel = torch.nonzero(~plain).flatten()
        if len(wsel):
            ow = owner[wsel]; pat = pattern[wsel]; j0 = jn_p[wsel]
            fc = first_code[ow]
            steps = W
            codes = st.walk(q0[ow], j0, ins_site0[ow], ins_off0[ow], lambda site: (pat >> (site - j0).clamp(min=0, max=KWIN)) & 1, steps)
            has_first = fc != 0
            if bool(has_first.any()):                    # the column's own character comes first, the walk supplies the other W - 1
                codes = torch.where(has_first[:, None], torch.cat([fc[:, None], codes[:, :W - 1]], dim=1), codes)
            h2, l2 = _pack_keys(codes); hi[wsel] = h2; lo[wsel] = l2
        dep = dep_site[owner] >= 0
        u = ~dep
        emit(hi[u], lo[u], col_id[owner][u], rank[u], prev_ch[owner][u], cnt[u], tag[owner][u])
        d = torch.nonzero(dep).flatten()
        if len(d):                                       # expand into haplotypes: bit i of the mask = haplotype of rank i in order[jn]
            md = m[d]; carr = st.PA[jn_p[d]][:, 0, :]
            bits = torch.arange(64, device=dev)
            mem = ((md[:, :, None] >> bits[None, None, :]) & 1).reshape(len(d), HW * 64)[:, :H].bool()
            car = ((carr[:, :, None] >> bits[None, None, :]) & 1).reshape(len(d), HW * 64)[:, :H]
            r, i = torch.nonzero(mem, as_tuple=True)
            site = dep_site[owner][d][r]
            kd = st.kind[site]; p = st.pos[site]; ln = st.len[site]
            c1 = torch.where(kd == SNP, st.alt[site].to(i64), torch.where(kd == DEL, st.ref[p.clamp(max=L)].to(i64), st.ins[(st.ins_off[site + 1] - 1).clamp(min=0)].to(i64)))
            c0 = torch.where(kd == DEL, st.ref[(p + ln).clamp(max=L)].to(i64), st.ref[p.clamp(max=L)].to(i64))
            chd = torch.where(car[r, i] == 1, c1, c0)
            dd = d[r]
            emit(hi[dd], lo[dd], col_id[owner][dd], i, chd, torch.ones_like(i), tag[owner][dd])

    neg1 = lambda n: torch.full((n,), -1, dtype=i64, device=dev)
    zero = lambda n: torch.zeros(n, dtype=i64, device=dev)
    # ---- backbone columns [0, END0), in chunks (bounded memory)
    CH = 1 << 24
    for a in range(0, END0, CH):
        c = torch.arange(a, min(END0, a + CH), device=dev); n = len(c)
        jn = torch.searchsorted(st.vs[:S], c, right=True)                   # first site with varstart > c
        jp = (jn - 1).clamp(min=0); has = jn > 0
        inside = has & (c < st.ve[jp])                                      # the reference allele of a SNP, or a deleted base: non-carriers only
        after = has & (c == st.ve[jp])                                      # right behind a site: the BWT symbol depends on its allele
        forced = torch.where(inside, zero(n), neg1(n))
        prev = torch.where(c > 0, st.ref[(c - 1).clamp(min=0)].to(i64), zero(n))
        rp = st.fwd_pos(c)
        classes(c, c, jn, forced, zero(n).to(torch.uint8), neg1(n), zero(n), prev, torch.where(after, jp, neg1(n)), st.tag_of_refpos(rp))
        if log: log("  strand %d: backbone columns %d / %d" % (st.model, min(END0, a + CH), L))
    # ---- alternative bases of SNPs and inserted bases
    sj = torch.arange(S, device=dev)
    snp = sj[st.kind[:S] == SNP]
    if len(snp):
        p = st.pos[snp]
        classes(L + snp, p + 1, snp + 1, torch.ones_like(snp), st.alt[snp], neg1(len(snp)), zero(len(snp)), st.ref[(p - 1).clamp(min=0)].to(i64), neg1(len(snp)),
                st.tag_of_refpos(st.fwd_pos(p)))
    insj = sj[st.kind[:S] == INS]
    if len(insj):
        ln = st.len[insj]; site = torch.repeat_interleave(insj, ln)
        off = torch.arange(len(site), device=dev) - torch.repeat_interleave(torch.cumsum(ln, 0) - ln, ln)
        p = st.pos[site]
        prev = torch.where(off == 0, st.ref[p].to(i64), st.ins[(st.ins_off[site] + off - 1).clamp(min=0)].to(i64))
        classes(L + S + st.ins_off[site] + off, p + 1, site + 1, torch.ones_like(site), zero(len(site)).to(torch.uint8), site, off, prev, neg1(len(site)),
                st.tag_of_refpos(st.anchor_fwd[site]))
    # ---- the last W + 64 columns: no sites here, every haplotype is its own class (its key runs into its own end marker)
    c = torch.arange(END0, L, device=dev).repeat_interleave(H); h = torch.arange(H, device=dev).repeat(L - END0)
    seq = 2 * h + st.model
    codes = st.walk(c, torch.full_like(c, S), neg1(len(c)), zero(len(c)), lambda site: torch.zeros_like(site), W, tail=seq)
    hi, lo = _pack_keys(codes)
    emit(hi, lo, c, h, st.ref[c - 1].to(i64), torch.ones_like(c), st.tag_of_refpos(st.fwd_pos(c)))
    # ---- the end markers themselves: suffix '$'_seq, BWT symbol = last base of the sequence, tag 0
    h = torch.arange(H, device=dev); seq = 2 * h + st.model
    codes = torch.zeros((H, W), dtype=torch.uint8, device=dev)
    for t in range(5): codes[:, 1 + t] = ((seq >> (3 * (4 - t))) & 7).to(torch.uint8)
    hi, lo = _pack_keys(codes)
    emit(hi, lo, torch.full_like(h, L + S + int(st.ins_off[S]) + 1), h, st.ref[L - 1].to(i64).expand(H), torch.ones_like(h), torch.zeros_like(h))
    return {k: torch.cat(v) for k, v in rec.items()}

# --------------------------------------------------------------------------------------------------------------- the whole collection
def prepare(ref_len, n_hap, seed=3, device="cpu", log=None, graph=None, **kw):
    """The graph and its two strands (haplotype orders at every site) on `device`: enough to sample reads."""
    import time
    t0 = time.time()
    dev = torch.device(device) if not isinstance(device, int) else torch.device("cuda", device)
    g = graph if graph is not None else make_graph(ref_len, n_hap, seed, **kw)
    if g.H > 512: raise ValueError("at most 512 haplotypes")
    gm = mirror(g)
    if log: log("pangraph: %d haplotypes x %d bp reference, %d sites (%.1fs)" % (g.H, g.L, g.S, time.time() - t0))
    strands = [Strand(g, 0, dev, g.L), Strand(gm, 1, dev, g.L, gm.orig_anchor)]
    if log: log("pangraph: haplotype orders at every site (%.1fs)" % (time.time() - t0))
    return g, strands

def build(ref_len, n_hap, seed=3, device="cpu", log=None, graph=None, prepared=None, host=True, **kw):
    """-> dict(rl_sym uint8, rl_len uint64, tag_val uint64, tag_len uint64 (numpy), n, n_seq, graph, strands, dev_runs)
    Sequences are hap0, rc(hap0), hap1, rc(hap1), ... each followed by its own end marker (written '\\n'), as in synth.py."""
    import time
    t0 = time.time()
    g, strands = prepared if prepared is not None else prepare(ref_len, n_hap, seed, device, log, graph, **kw)
    dev = strands[0].dev
    recs = [strand_records(s, log) for s in strands]
    R = {k: torch.cat([r[k] for r in recs]) for k in recs[0]}
    del recs
    n_rec = len(R["hi"])
    if log: log("pangraph: %d records, sorting (%.1fs)" % (n_rec, time.time() - t0))
    return _assemble(R, strands, g, dev, log, t0, host)

def _assemble(R, strands, g, dev, log, t0, host=True):
    import time
    # stable LSD sort: sub, then lo, then hi
    perm = torch.argsort(R["sub"], stable=True)
    perm = perm[torch.argsort(R["lo"][perm], stable=True)]
    perm = perm[torch.argsort(R["hi"][perm], stable=True)]
    hi, lo, sub = R["hi"][perm], R["lo"][perm], R["sub"][perm]
    same_key = (hi[1:] == hi[:-1]) & (lo[1:] == lo[:-1])
    clash = same_key & ((sub[1:] >> 10) != (sub[:-1] >> 10))
    if bool(clash.any()):
        raise RuntimeError("pangraph: %d pairs of different columns share their first %d characters; the order is not decided (reseed)" % (int(clash.sum()), W))
    ch = R["ch"][perm]; cnt = R["cnt"][perm].to(torch.int64); tag = R["tag"][perm]
    def rle(val):
        start = torch.ones(len(val), dtype=torch.bool, device=dev); start[1:] = val[1:] != val[:-1]
        idx = torch.nonzero(start).flatten()
        csum = torch.cumsum(cnt, 0)
        ends = torch.cat([csum[idx[1:] - 1], csum[-1:]])
        lens = ends - torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), ends[:-1]])
        return val[idx], lens
    sym, slen = rle(ch.to(torch.int64))
    tv, tl = rle(tag)
    n = int(cnt.sum())
    if log: log("pangraph: n = %d, r = %d BWT runs, %d tag runs (%.1fs)" % (n, len(sym), len(tv), time.time() - t0))
    byte_of = torch.as_tensor(BYTE_OF, device=dev)
    dev_runs = dict(sym=byte_of[sym].contiguous(), len=slen.contiguous(), tag_val=tv.contiguous(), tag_len=tl.contiguous())
    out = dict(n=n, n_seq=2 * g.H, graph=g, strands=strands, dev_runs=dev_runs, r=len(sym), n_tag_runs=len(tv))
    if host: out.update(host_runs(dev_runs))            # numpy copies (7 GB for config 4): only where the CPU tools need them
    return out

def host_runs(dev_runs):
    """numpy copies of the run arrays (what prep.rlbwt_from_runs / tags_from_runs take)."""
    return dict(rl_sym=dev_runs["sym"].cpu().numpy(), rl_len=dev_runs["len"].cpu().numpy().astype(np.uint64),
                tag_val=dev_runs["tag_val"].cpu().numpy().astype(np.uint64), tag_len=dev_runs["tag_len"].cpu().numpy().astype(np.uint64))

# --------------------------------------------------------------------------------------------------------------- reads
def sample_reads(strands, n_reads, read_len, sub, seed, indel=0.0):
    """Reads off the haplotypes: uniform strand, haplotype and backbone start column (the haplotype's own path from there,
    through its alleles), i.i.d. substitutions (a substitution always changes the base) and, for the long-read config,
    1-bp insertions / deletions at rate `indel`. -> uint8 [n_reads, read_len] on the strands' device."""
    dev = strands[0].dev; g = torch.Generator(device=dev); g.manual_seed(seed)
    margin = read_len // 8 + 16 if indel > 0 else 0
    span = read_len + margin
    out = torch.empty((n_reads, read_len), dtype=torch.uint8, device=dev)
    byte_of = torch.as_tensor(BYTE_OF, device=dev)
    acgt = torch.tensor([A, C, G, T], dtype=torch.uint8, device=dev)
    model = torch.randint(0, 2, (n_reads,), generator=g, device=dev)
    step = max(1, (1 << 27) // span)
    for st in strands:
        if not hasattr(st, "alleles_dev"):
            st.alleles_dev = torch.as_tensor(np.ascontiguousarray(st.g.alleles.T), device=dev)       # [S, H]
        sel = torch.nonzero(model == st.model).flatten()
        for a in range(0, len(sel), step):
            ids = sel[a:a + step]; m = len(ids)
            h = torch.randint(0, st.H, (m,), generator=g, device=dev)
            c = torch.randint(0, st.L - span - 12 * KWIN - 64, (m,), generator=g, device=dev)
            jn = torch.searchsorted(st.vs[:st.S], c, right=False)                                     # first site with varstart >= c
            jp = (jn - 1).clamp(min=0)
            inside = (jn > 0) & (c < st.ve[jp])                                                       # inside a deleted stretch: start behind it
            c = torch.where(inside, st.ve[jp], c)
            S = st.S
            bit_of = lambda site: st.alleles_dev[site.clamp(max=S - 1), h].to(torch.int64) * (site < S)
            codes = st.walk(c, jn, torch.full_like(c, -1), torch.zeros_like(c), bit_of, span)
            r = byte_of[codes.long()]
            if indel > 0:
                ev = torch.rand((m, read_len), generator=g, device=dev)
                dele = (ev < indel / 2).to(torch.int64); inse = (ev >= indel / 2) & (ev < indel)
                srcoff = (torch.cumsum(1 + dele - inse.to(torch.int64), 1) - 1).clamp(min=0, max=span - 1)
                r = torch.gather(r, 1, srcoff)
                r = torch.where(inse, acgt[torch.randint(0, 4, r.shape, generator=g, device=dev)], r)
            else:
                r = r[:, :read_len]
            mm = torch.rand(r.shape, generator=g, device=dev) < sub
            ridx = torch.randint(0, 4, r.shape, generator=g, device=dev)
            alt = torch.where(acgt[ridx] == r, acgt[(ridx + 1) % 4], acgt[ridx])
            out[ids] = torch.where(mm, alt, r)
    return out
