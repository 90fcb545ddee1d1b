#!/usr/bin/env python3
"""Tuning run (GPU box): index-layout knobs and K1 variants against fmd_mem_kernel time on the bench workload. One process, one
workload build (the run arrays stay on the GPU), one index per setting (pgi_index_create_from_runs, ~0.3 s).
    python tools/index_sweep.py --settings "default;PGI_BSHIFT=10;PGI_SEEDK=12;PGI_K1=15;PGI_BSHIFT=10,PGI_K1=14" [bench.py flags]
Every setting is a comma-separated list of NAME=value environment assignments (cleared again afterwards)."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench

def main():
    settings = "default"
    if "--settings" in sys.argv:
        i = sys.argv.index("--settings"); settings = sys.argv[i + 1]; del sys.argv[i:i + 2]
    args = bench.parse_args()
    import torch
    from __graft_entry__ import load_package
    capi = load_package().capi
    torch.cuda.set_device(0)
    w = bench.make_workload(args, 0, 0, need_index=False)
    runs = w.runs
    L = args.read_len; n = args.reads_per_step; pool = w.reads.shape[0]
    d_off = torch.arange(n + 1, dtype=torch.int64, device="cuda") * L
    d_reads = [torch.cat([w.reads[i].reshape(-1), torch.zeros(16, dtype=torch.uint8, device="cuda")]) for i in range(pool)]
    del w.reads
    out = {}; ref = None
    for setting in settings.split(";"):
        env = {} if setting == "default" else dict(kv.split("=") for kv in setting.split(","))
        for k, v in env.items(): os.environ[k] = v
        t0 = time.time()
        ix = capi.Index.from_runs(runs["sym"], runs["len"], runs["tag_val"], runs["tag_len"], device=0)
        torch.cuda.synchronize(); t_load = time.time() - t0
        batch = capi.Batch(ix)
        ms = []; tot = []; sig = None
        for i in range(args.warmup + args.steps):
            batch.set_device_reads(d_reads[i % pool].data_ptr(), d_off.data_ptr(), n, n * L)
            batch.run(args.min_len, args.min_occ)
            if i >= args.warmup: st = batch.stage_ms(); ms.append(st[0]); tot.append(st[5])
            if i == 0:
                r = batch.device_result(); sig = (int(r.n_mems), int(r.n_tags), int(r.extensions)); blk = int(r.blocks_fetched); sl = int(r.seed_lookups)
        qt = []
        for i in range(args.warmup + args.steps):
            batch.set_device_reads(d_reads[i % pool].data_ptr(), d_off.data_ptr(), n, n * L)
            batch.run_query_tags()
            if i >= args.warmup: qt.append(batch.stage_ms()[5])
        if ref is None: ref = sig
        out[setting] = {"k1_ms_median": round(float(np.median(ms)), 3), "k1_ms_min": round(float(min(ms)), 3), "step_ms_median": round(float(np.median(tot)), 3), "query_tags_step_ms": round(float(np.median(qt)), 3), "same_counts_as_first": sig == ref,
                        "blocks_fetched": blk, "seed_lookups": sl, "bucket_shift": ix.info["bucket_shift"], "seed_k": ix.info["seed_k"], "n_dev_blocks": ix.info["n_dev_blocks"],
                        "n_crowded_buckets": ix.info["n_crowded_buckets"], "device_bytes": ix.info["device_bytes"], "index_create_s": round(t_load, 2)}
        print("[sweep]", setting, out[setting], file=sys.stderr, flush=True)
        batch.close(); ix.close()
        for k in env: os.environ.pop(k, None)
    print(json.dumps({"reads_per_step": n, "facts": w.facts, "settings": out}))

if __name__ == "__main__":
    main()
