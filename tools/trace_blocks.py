#!/usr/bin/env python3
"""CPU analysis (no GPU): which run blocks does fmd_mem_kernel touch? Builds a small pangenome of the bench's kind, runs
the product's own lane state machine (tests/hostsim) over reads sampled from it with the extension trace on, and reports
per executed extension: positions in one bucket, in one 128-byte line, crowded-bucket hits by phase, the hottest buckets.
    python tools/trace_blocks.py [ref_len n_hap n_reads [bshift [seed_k]]]"""
import ctypes, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_package
import sim

def main():
    ref_len = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    n_hap = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    n_reads = int(sys.argv[3]) if len(sys.argv) > 3 else 400
    if len(sys.argv) > 4 and int(sys.argv[4]): os.environ["PGI_BSHIFT"] = sys.argv[4]
    seed_k = int(sys.argv[5]) if len(sys.argv) > 5 else 10
    pkg = load_package(); pg = pkg.pangraph; prep = pkg.prep
    prepared = pg.prepare(ref_len, n_hap, seed=3, device="cpu")
    d = pg.build(ref_len, n_hap, prepared=prepared, device="cpu")
    ri, info = prep.rindex_from_rlbwt(prep.rlbwt_from_runs(np.asarray(d["rl_sym"]), np.asarray(d["rl_len"])), encoded=True, with_samples=False)
    g = sim.Index(ri); L = sim.lib()
    s = g.info["bshift"]; nb = g.info["n_buckets"] if "n_buckets" in g.info else (g.info["n"] >> s) + 1
    print("n = %d, r = %d, bucket shift %d, blocks %d, seed K %d" % (g.info["n"], len(d["rl_sym"]), s, g.info["n_blocks"], g.set_seed_k(seed_k)), file=sys.stderr)
    reads = pg.sample_reads(prepared[1], n_reads, 150, 0.01, 5)
    reads = reads.cpu().numpy() if hasattr(reads, "cpu") else np.asarray(reads)
    L.sim_trace.argtypes = [ctypes.c_int]; L.sim_trace_read.restype = ctypes.c_uint64; L.sim_trace_read.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
    L.sim_block.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
    L.sim_trace(1)
    tot_ext = 0
    for i in range(n_reads):
        _, ne = g.find_mems(reads[i].tobytes(), 20, 1); tot_ext += ne
    cnt = L.sim_trace_read(None, 0); buf = np.zeros(cnt, dtype=np.uint64); L.sim_trace_read(buf.ctypes.data, cnt); L.sim_trace(0)
    t = buf.reshape(-1, 3); n = g.info["n"]
    p1 = np.minimum(t[:, 0], n); p2 = np.minimum(t[:, 1], n); ph = t[:, 2]
    b1 = p1 >> np.uint64(s); b2 = p2 >> np.uint64(s)
    mask = np.uint64((1 << s) - 1)
    # directory flag of every touched bucket
    ub = np.unique(np.concatenate([b1, b2])); isdir = {}
    w = np.zeros(16, dtype=np.uint32)
    for b in ub:
        L.sim_block(g.h, int(b), w.ctypes.data); isdir[int(b)] = (int(w[9]) & 0xF) == 0xF
    d1 = np.array([isdir[int(b)] for b in b1]) & ((p1 & mask) != 0)
    d2 = np.array([isdir[int(b)] for b in b2]) & ((p2 & mask) != 0)
    E = len(t)
    out = {"reads": n_reads, "reference_extensions": int(tot_ext), "executed_extensions": E, "bucket_shift": int(s),
           "same_bucket": float((b1 == b2).mean()), "same_128B_line": float(((b1 >> np.uint64(1)) == (b2 >> np.uint64(1))).mean()),
           "chunk_fetches_per_extension": float((d1.sum() + d2.sum()) / E),
           "extensions_with_a_chunk": float((d1 | d2).mean()),
           "interval_size_quantiles": [int(x) for x in np.quantile((p2 - p1).astype(np.float64), [0.1, 0.5, 0.9])]}
    for name, code in (("P1", 1), ("P2", 2), ("P3", 3)):
        m = ph == code
        out["phase_" + name] = {"share": float(m.mean()), "chunk_per_ext": float((d1[m].sum() + d2[m].sum()) / max(1, m.sum())), "same_bucket": float((b1[m] == b2[m]).mean())}
    hot = {}
    for bb, dd in ((b1, d1), (b2, d2)):
        u, c = np.unique(bb[dd], return_counts=True)
        for x, y in zip(u, c): hot[int(x)] = hot.get(int(x), 0) + int(y)
    top = sorted(hot.items(), key=lambda x: -x[1])[:12]
    out["hottest_crowded_buckets"] = [{"bucket": b, "first_position": b << int(s), "hits": c} for b, c in top]
    out["crowded_buckets_touched"] = len(hot); out["C"] = [int(x) for x in np.cumsum([0] + [int((np.asarray(d["rl_len"])[np.asarray(d["rl_sym"]) == c]).sum()) for c in (10, 65, 67, 71, 78, 84)])]
    print(json.dumps(out, indent=1))

if __name__ == "__main__":
    main()
