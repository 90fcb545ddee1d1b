#!/usr/bin/env python3
"""Timeline of one pipelined pgi_find_mems call (PGI_TRACE=1) on a small synthetic workload. GPU box only."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from __graft_entry__ import load_package
pkg = load_package(); capi = pkg.capi; prep = pkg.prep
from pangenome_index_b200 import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000000
d = synth.build(4_000_000, 8, seed=3, device=0)
reads = synth.sample_reads(d["text"], d["seq_starts"], n, 150, 0.01, 5, torch.device("cuda", 0))
ri, info = prep.rindex_from_rlbwt(prep.rlbwt_from_runs(d["rl_sym"], d["rl_len"]), encoded=True, with_samples=False)
tags, _ = prep.tags_from_runs(d["tag_val"], d["tag_len"], 0)
ix = capi.Index(ri, tags, 0)
h = torch.empty(n * 150 + 16, dtype=torch.uint8).pin_memory(); h[: n * 150].copy_(reads.reshape(-1))
off = (torch.arange(n + 1, dtype=torch.int64) * 150).pin_memory()
for i in range(4):
    t = time.perf_counter(); r = ix.find_mems_raw(h.data_ptr(), off.data_ptr(), n, 20, 1); dt = time.perf_counter() - t
    print("call %d: %.2f ms  mems %d tags %d" % (i, dt * 1e3, r.n_mems, r.n_tags), flush=True); ix.result_free(r)
import threading
h2 = torch.empty(n * 150 + 16, dtype=torch.uint8).pin_memory(); h2.copy_(h)
def loop(buf, k):
    for _ in range(k):
        r = ix.find_mems_raw(buf.data_ptr(), off.data_ptr(), n, 20, 1); ix.result_free(r)
for nth in (1, 3):
    bufs = [h, h2, h][:nth]
    for rep in range(2):
        ths = [threading.Thread(target=loop, args=(bufs[i], 4)) for i in range(nth)]
        t = time.perf_counter(); [x.start() for x in ths]; [x.join() for x in ths]; dt = time.perf_counter() - t
    print("%d concurrent callers x 4 calls: %.2f ms per call, %.1f M reads/s" % (nth, dt * 1e3 / (4 * nth), 4 * nth * n / dt / 1e6), flush=True)
# raw copy bandwidths for reference
a = torch.empty(300_000_000, dtype=torch.uint8, device="cuda"); hb = torch.empty(300_000_000, dtype=torch.uint8).pin_memory()
for name, f in (("H2D", lambda: a.copy_(hb, non_blocking=True)), ("D2H", lambda: hb.copy_(a, non_blocking=True))):
    f(); torch.cuda.synchronize(); t = time.perf_counter(); f(); torch.cuda.synchronize(); print(name, "%.1f GB/s" % (0.3 / (time.perf_counter() - t)))
