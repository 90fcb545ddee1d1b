#!/usr/bin/env python3
"""How large a synthetic pangenome the in-bench GPU builder can make: time and peak memory. GPU box only."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from __graft_entry__ import load_package
pkg = load_package(); prep = pkg.prep
from pangenome_index_b200 import synth
ref_len, n_hap = int(sys.argv[1]), int(sys.argv[2])
t = time.time(); torch.cuda.reset_peak_memory_stats()
d = synth.build(ref_len, n_hap, seed=3, device=0, log=lambda *a: print(*a, flush=True))
print("build %.1fs peak %.1f GB n=%d r=%d tag_runs=%d" % (time.time() - t, torch.cuda.max_memory_allocated() / 1e9, d["n"], len(d["rl_sym"]), len(d["tag_val"])), flush=True)
t = time.time(); rl = prep.rlbwt_from_runs(d["rl_sym"], d["rl_len"]); ri, info = prep.rindex_from_rlbwt(rl, encoded=True, with_samples=False)
print("rindex %.1fs %d bytes" % (time.time() - t, len(ri)), info, flush=True)
t = time.time(); tags, nr = prep.tags_from_runs(d["tag_val"], d["tag_len"], 0); print("tags %.1fs %d bytes %d runs" % (time.time() - t, len(tags), nr), flush=True)
t = time.time(); ix = pkg.capi.Index(ri, tags, 0); print("index %.1fs" % (time.time() - t), ix.info, flush=True)
