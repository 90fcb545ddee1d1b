#!/usr/bin/env python3
"""Build a synthetic pangenome index at a given scale on the GPU and report the cost of every step (data prep + GPU-side
index construction). Usage: python tools/pangraph_scale.py REF_LEN N_HAP [--reads N]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from __graft_entry__ import load_package
def main():
    ref_len, n_hap = int(float(sys.argv[1])), int(sys.argv[2]); n_reads = int(float(sys.argv[4])) if len(sys.argv) > 4 else 200000
    pkg = load_package(); pg = pkg.pangraph; capi = pkg.capi
    out = {"ref_len": ref_len, "n_hap": n_hap}
    log = lambda *a: print("[scale]", *a, file=sys.stderr, flush=True)
    torch.cuda.reset_peak_memory_stats(); t0 = time.time()
    d = pg.build(ref_len, n_hap, seed=3, device=0, log=log)
    torch.cuda.synchronize(); out["pangraph_s"] = round(time.time() - t0, 2); out["pangraph_peak_gb"] = round(torch.cuda.max_memory_allocated() / 1e9, 2)
    out.update(n=d["n"], r=len(d["rl_sym"]), tag_runs=len(d["tag_val"]), sites=d["graph"].S)
    dr = d["dev_runs"]; t0 = time.time()
    ix = capi.Index.from_runs(dr["sym"].contiguous(), dr["len"].contiguous(), dr["tag_val"].contiguous(), dr["tag_len"].contiguous(), device=0)
    out["from_runs_s"] = round(time.time() - t0, 2); out["index"] = {k: ix.info[k] for k in ("n", "n_dev_blocks", "bucket_shift", "n_crowded_buckets", "device_bytes", "tag_items", "tag_item_width", "seed_k", "n_runs")}
    t0 = time.time(); reads = pg.sample_reads(d["strands"], n_reads, 150, 0.01, seed=4); torch.cuda.synchronize(); out["sample_reads_s"] = round(time.time() - t0, 2)
    r = reads.reshape(-1).cpu().numpy(); off = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(150)
    t0 = time.time(); G = ix.find_mems(r, off, 20, 1, text=False); out["find_mems_s"] = round(time.time() - t0, 3)
    out.update(mems_per_read=G["n_mems"] / n_reads, tags_per_read=G["n_tags"] / n_reads, ext_per_read=G["extensions"] / n_reads,
               full_length_reads=float(((G["mems"][:, 0] == 0) & (G["mems"][:, 1] == 150)).sum()) / n_reads, max_bwt_start=int(G["mems"][:, 2].max()))
    if "--oracle" in sys.argv:
        import orc
        prep = pkg.prep; t0 = time.time()
        ri, info = prep.rindex_from_rlbwt(prep.rlbwt_from_runs(d["rl_sym"], d["rl_len"]), encoded=True, with_samples=False)
        tags, _ = prep.tags_from_runs(d["tag_val"], d["tag_len"], 0); out["prep_files_s"] = round(time.time() - t0, 1); out["ri_bytes"] = len(ri); out["tags_bytes"] = len(tags)
        t0 = time.time(); o = orc.Index(ri, tags, 0); out["oracle_load_s"] = round(time.time() - t0, 1)
        ns = min(n_reads, 20000)
        R = o.find_mems(r[: ns * 150], off[: ns + 1], 20, 1, text=False, count_work=True); G2 = ix.find_mems(r[: ns * 150], off[: ns + 1], 20, 1, text=False)
        out["parity"] = bool(np.array_equal(R["mems"], G2["mems"]) and np.array_equal(R["tags"], G2["tags"]) and np.array_equal(R["n_runs"], G2["n_runs"]) and G2["extensions"] == R["counters"]["E"])
        out["oracle_reads_per_s"] = ns / R["seconds"]
    print(json.dumps(out))
if __name__ == "__main__":
    main()
