#!/usr/bin/env python3
"""Runs the REFERENCE'S OWN find_mems code (oracle/_ref/libpgiref.so: src/r-index.cpp + src/tag_arrays.cpp unmodified,
algorithm.hpp:644-757 and the main loops of src/find_mems.cpp / src/query_tags.cpp cut out by line range; built by
oracle/Makefile against oracle/ref_shim) on the cases of tests/refcases.py and records what it printed / returned in
tests/golden/ref_find_mems.json. Runs only where /root/reference is mounted (the build container); the GPU box and any
other machine compare against the committed file.

Usage: python tests/golden/make_ref_golden.py
"""
import hashlib, json, os, sys
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import numpy as np
import orc, refcases

def sha(b): return hashlib.sha256(b).hexdigest()

def extend_vectors(n, seed, N=20000):
    """Seeded random (interval, symbol, direction) triples inside [0, n] and rank positions (no undefined inputs:
    the reference reads out of bounds for positions beyond n and for bytes >= 0x80)."""
    rng = np.random.default_rng(seed)
    iv = np.zeros((N, 3), dtype=np.uint64)
    k = rng.integers(0, n, N); sz = np.minimum(rng.integers(0, 300, N), n - k)
    iv[:, 0] = k; iv[:, 1] = rng.integers(0, n, N); iv[:, 2] = sz
    iv[:200] = [0, 0, n]
    sym = rng.choice(np.frombuffer(b"ACGTN\n\0acgtnX", dtype=np.uint8), N); d = rng.integers(0, 2, N).astype(np.uint8)
    pos = np.concatenate([rng.integers(0, n + 1, 5000), [0, n, n - 1, 1]]).astype(np.uint64)
    return iv, sym, d, pos

def main():
    if not orc.ref_available():
        sys.exit("oracle/_ref/libpgiref.so cannot be built here (no /root/reference)")
    IX = refcases.indexes(); out = {"_about": __doc__.strip().split("\n\n")[0], "find_mems": {}, "extend_rank": {}, "query_tags": {}}
    total_reads = 0
    for c in refcases.CASES:
        ix = IX[c[0]]; R = orc.Index(ix["ri"], ix["tags"], ix["layout"], impl="ref")
        rs = refcases.reads_of(ix, c[1]); total_reads += len(rs)
        text = R.ref_text(b"\n".join(rs) + b"\n", c[2], c[3], verbatim=True)
        e = {"n_reads": len(rs), "stdout_sha256": sha(text), "stdout_bytes": len(text), "n_mem_lines": text.count(b"MEM START"),
             "verbatim_main_loop": bool(ix["layout"] == 0 and ix["tags"] is not None)}
        if len(text) < 3000: e["stdout"] = text.decode()
        out["find_mems"][refcases.case_key(c)] = e
        print(refcases.case_key(c), e["n_reads"], e["n_mem_lines"], e["stdout_bytes"])
    for name, ix in IX.items():
        R = orc.Index(ix["ri"], None, 0, impl="ref")
        iv, sym, d, pos = extend_vectors(R.n, 3)
        out["extend_rank"][name] = {"n": R.n, "extend_sha256": sha(R.extend_batch(iv, sym, d).tobytes()), "rank_sha256": sha(R.rank_batch(pos).tobytes())}
    for name in ("synth_enc_compact", "synth_leg_compact", "xy_legacy_compact"):
        ix = IX[name]; R = orc.Index(ix["ri"], ix["tags"], 0, impl="ref")
        rs = refcases.reads_of(ix, ("sample", 400, 40, 0.01, 12))
        text = R.ref_query_tags_text(b"\n".join(rs) + b"\n")
        out["query_tags"][name] = {"n_reads": len(rs), "stdout_sha256": sha(text), "stdout_bytes": len(text)}
    out["_reads_total"] = total_reads
    json.dump(out, open(os.path.join(HERE, "ref_find_mems.json"), "w"), indent=1, sort_keys=True)
    print("reference outputs recorded for", total_reads, "reads")
if __name__ == "__main__":
    main()
