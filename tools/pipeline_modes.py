#!/usr/bin/env python3
"""Tuning run (GPU box): how steps share the GPU. One index build, then
  (1) device-resident steps with 1 / 2 / 3 batches in flight (pgi_batch_run_async / pgi_batch_wait) for PGI_CHAIN = 1, 3, 2,
  (2) end-to-end pgi_find_mems_compact calls from 3 / 4 host threads for PGI_CHAIN = 0, 1, 3, 2 (K = --steps calls, as bench.py
      times them, and 3 K calls: the pipeline's fill and drain amortised).
Prints one JSON object.   python tools/pipeline_modes.py [bench.py flags]"""
import json, os, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench

def main():
    args = bench.parse_args()
    import torch
    from __graft_entry__ import load_package
    capi = load_package().capi
    torch.cuda.set_device(0)
    w = bench.make_workload(args, 0, 0)
    ix = w.ix
    L = args.read_len; n = args.reads_per_step; pool = w.reads.shape[0]
    d_off = torch.arange(n + 1, dtype=torch.int64, device="cuda") * L
    h_reads = [torch.empty(n * L + 16, dtype=torch.uint8).pin_memory() for _ in range(pool)]
    for i in range(pool): h_reads[i][: n * L].copy_(w.reads[i].reshape(-1))
    d_reads = [torch.cat([w.reads[i].reshape(-1), torch.zeros(16, dtype=torch.uint8, device="cuda")]) for i in range(pool)]
    del w.reads
    torch.cuda.synchronize()
    out = {"reads_per_step": n, "facts": w.facts, "device": {}, "e2e": {}}
    K = args.steps
    # ---- (1) device-resident, `depth` batches in flight from one host thread
    batches = [capi.Batch(ix) for _ in range(3)]
    ref = None
    for mode in ("1", "3", "2"):
        os.environ["PGI_CHAIN"] = mode
        for depth in (1, 2, 3):
            def run(k_steps):
                inflight = []
                for i in range(k_steps):
                    b = batches[i % depth]
                    if len(inflight) == depth: inflight.pop(0).wait()
                    b.set_device_reads(d_reads[i % pool].data_ptr(), d_off.data_ptr(), n, n * L)
                    b.run_async(args.min_len, args.min_occ); inflight.append(b)
                for b in inflight: b.wait()
            run(max(args.warmup, depth)); torch.cuda.synchronize()
            t0 = time.perf_counter(); run(K); torch.cuda.synchronize(); ms = 1e3 * (time.perf_counter() - t0) / K
            r = batches[(K - 1) % depth].device_result(); sig = (int(r.n_mems), int(r.n_tags), int(r.extensions))
            k1 = [batches[j].stage_ms()[0] for j in range(depth)]
            if ref is None: ref = sig
            out["device"]["chain%s_depth%d" % (mode, depth)] = {"ms_per_step": round(ms, 3), "Mreads_s": round(n / ms / 1e3, 1), "k1_ms_last": [round(x, 3) for x in k1], "same_counts": sig == ref}
            print("[device] chain", mode, "depth", depth, out["device"]["chain%s_depth%d" % (mode, depth)], file=sys.stderr, flush=True)
    for b in batches: b.close()
    # ---- (2) end to end, compact wire form, concurrent callers
    def e2e_pass(k_calls, callers):
        nxt = [0]; lock = threading.Lock()
        def caller():
            while True:
                with lock:
                    i = nxt[0]
                    if i >= k_calls: return
                    nxt[0] += 1
                r = ix.find_mems_compact_raw(h_reads[i % pool].data_ptr(), None, L, n, args.min_len, args.min_occ)
                ix.result_compact_free(r)
        ths = [threading.Thread(target=caller) for _ in range(min(callers, k_calls))]
        t0 = time.perf_counter(); [t.start() for t in ths]; [t.join() for t in ths]
        return 1e3 * (time.perf_counter() - t0) / k_calls
    for mode in ("0", "1", "3", "2"):
        os.environ["PGI_CHAIN"] = mode
        for callers in (3, 4):
            e2e_pass(max(args.warmup, callers), callers)
            a = min(e2e_pass(K, callers) for _ in range(2)); b3 = e2e_pass(3 * K, callers)
            out["e2e"]["chain%s_callers%d" % (mode, callers)] = {"ms_per_call_K": round(a, 3), "Mreads_s_K": round(n / a / 1e3, 1), "ms_per_call_3K": round(b3, 3), "Mreads_s_3K": round(n / b3 / 1e3, 1)}
            print("[e2e] chain", mode, "callers", callers, out["e2e"]["chain%s_callers%d" % (mode, callers)], file=sys.stderr, flush=True)
    print(json.dumps(out))

if __name__ == "__main__":
    main()
