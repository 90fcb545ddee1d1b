#!/usr/bin/env python3
"""Tuning run: bucket size of the run-block array (PGI_BSHIFT = log2 positions per 64-byte block; default: chosen by the
flattener) against fmd_mem_kernel time on the bench workload. One process, one workload build, one index per value.
    python tools/bshift_sweep.py --shifts auto,6,7,8 [bench.py flags]"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench

def main():
    shifts = "auto,6,7,8"
    if "--shifts" in sys.argv:
        i = sys.argv.index("--shifts"); shifts = sys.argv[i + 1]; del sys.argv[i:i + 2]
    args = bench.parse_args()
    import torch
    from __graft_entry__ import load_package
    capi = load_package().capi
    torch.cuda.set_device(0)
    w = bench.make_workload(args, 0, 0)
    L = args.read_len; n = args.reads_per_step; pool = w["reads"].shape[0]
    d_off = torch.arange(n + 1, dtype=torch.int64, device="cuda") * L
    stream = torch.cuda.current_stream()
    d_reads = [torch.cat([w["reads"][i].reshape(-1), torch.zeros(16, dtype=torch.uint8, device="cuda")]) for i in range(pool)]
    out = {}; ref = None
    for v in shifts.split(","):
        if v == "auto": os.environ.pop("PGI_BSHIFT", None)
        else: os.environ["PGI_BSHIFT"] = v
        t0 = time.time()
        ix = capi.Index(w["ri"], w["tags"], w["layout"], device=0)
        t_load = time.time() - t0
        batch = capi.Batch(ix, stream=stream.cuda_stream)
        ms = []; sig = None
        for i in range(args.warmup + args.steps):
            batch.set_device_reads(d_reads[i % pool].data_ptr(), d_off.data_ptr(), n, n * L)
            batch.run(args.min_len, args.min_occ)
            if i >= args.warmup: ms.append(batch.stage_ms()[0])
            if i == 0:
                r = batch.device_result(); sig = (int(r.n_mems), int(r.n_tags), int(r.extensions))
        if ref is None: ref = sig
        out[v] = {"k1_ms_median": float(np.median(ms)), "k1_ms_min": float(min(ms)), "counts": sig, "same_counts_as_first": sig == ref,
                  "bucket_shift": ix.info["bucket_shift"], "n_dev_blocks": ix.info["n_dev_blocks"], "n_crowded_buckets": ix.info["n_crowded_buckets"],
                  "device_bytes": ix.info["device_bytes"], "index_create_s": round(t_load, 2)}
        print("[sweep]", v, out[v], file=sys.stderr, flush=True)
        batch.close(); ix.close()
    print(json.dumps({"reads_per_step": n, "facts": w["facts"], "shifts": out}))

if __name__ == "__main__":
    main()
