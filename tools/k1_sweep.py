#!/usr/bin/env python3
"""Tuning run: time fmd_mem_kernel variants (PGI_K1 = "<fetch mode><CTAs per SM>") on the bench workload, one process,
one index build. Prints per-variant kernel ms (median of --steps device-resident steps) and checks that every variant
returns the same MEMs / tags as the first one.   python tools/k1_sweep.py --variants 04,14,15,24 [bench.py flags]"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench

def main():
    variants = "14,24"
    if "--variants" in sys.argv:
        i = sys.argv.index("--variants"); variants = sys.argv[i + 1]; del sys.argv[i:i + 2]
    args = bench.parse_args()
    import torch
    from __graft_entry__ import load_package
    capi = load_package().capi
    torch.cuda.set_device(0)
    w = bench.make_workload(args, 0, 0)
    ix = w.ix
    L = args.read_len; n = args.reads_per_step; pool = w.reads.shape[0]
    d_off = torch.arange(n + 1, dtype=torch.int64, device="cuda") * L
    stream = torch.cuda.current_stream()
    batch = capi.Batch(ix, stream=stream.cuda_stream)
    d_reads = [torch.cat([w.reads[i].reshape(-1), torch.zeros(16, dtype=torch.uint8, device="cuda")]) for i in range(pool)]
    out = {}; ref = None
    for v in variants.split(","):
        os.environ["PGI_K1"] = v
        ms = []; sig = None
        for i in range(args.warmup + args.steps):
            batch.set_device_reads(d_reads[i % pool].data_ptr(), d_off.data_ptr(), n, n * L)
            batch.run(args.min_len, args.min_occ)
            if i >= args.warmup: ms.append(batch.stage_ms()[0])
            if i == 0:
                r = batch.device_result(); sig = (int(r.n_mems), int(r.n_tags), int(r.extensions))
        if ref is None: ref = sig
        out[v] = {"k1_ms_median": float(np.median(ms)), "k1_ms_min": float(min(ms)), "counts": sig, "same_counts_as_first": sig == ref}
        print("[sweep]", v, out[v], file=sys.stderr, flush=True)
    print(json.dumps({"reads_per_step": n, "facts": w.facts, "variants": out}))

if __name__ == "__main__":
    main()
