#!/usr/bin/env python3
"""End-to-end rate of the find_mems CLI (host/find_mems: reads file in, the reference's stdout text out to /dev/null) on a
synthetic pangenome: what a user of the reference's command line sees. Usage: python tools/cli_e2e.py [REF_LEN N_HAP N_READS]"""
import json, os, re, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from __graft_entry__ import load_package
def main():
    ref_len = int(float(sys.argv[1])) if len(sys.argv) > 1 else 50_000_000; n_hap = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    n_reads = int(float(sys.argv[3])) if len(sys.argv) > 3 else 8_000_000
    pkg = load_package(); pg = pkg.pangraph; prep = pkg.prep
    d = pg.build(ref_len, n_hap, seed=3, device=0)
    ri, _ = prep.rindex_from_rlbwt(prep.rlbwt_from_runs(d["rl_sym"], d["rl_len"]), encoded=True, with_samples=False)
    tags, _ = prep.tags_from_runs(d["tag_val"], d["tag_len"], 0)
    tmp = tempfile.mkdtemp(prefix="pgi_cli_")
    open(os.path.join(tmp, "p.ri"), "wb").write(ri); open(os.path.join(tmp, "p.tags"), "wb").write(tags)
    with open(os.path.join(tmp, "reads.txt"), "wb") as f:
        for a in range(0, n_reads, 2_000_000):
            r = pg.sample_reads(d["strands"], min(2_000_000, n_reads - a), 150, 0.01, seed=11 + a)
            nl = torch.full((r.shape[0], 1), 10, dtype=torch.uint8, device=r.device)
            f.write(torch.cat([r, nl], dim=1).cpu().numpy().tobytes())
    del d; torch.cuda.empty_cache()
    out = {"n": None, "reads": n_reads, "ri_bytes": len(ri), "tags_bytes": len(tags)}
    exe = os.path.join(ROOT, "pangenome-index_b200", "host", "find_mems")
    for label, extra in (("text_to_dev_null", []), ("no_text", ["--no-text"])):
        t0 = time.time()
        p = subprocess.run([exe, os.path.join(tmp, "p.ri"), os.path.join(tmp, "p.tags"), os.path.join(tmp, "reads.txt"), "20", "1"] + extra, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True)
        m = re.search(r"in ([0-9.e+-]+) s \(([0-9.e+-]+) reads/s", p.stderr); ld = re.search(r"flattening the index took ([0-9.e+-]+)", p.stderr)
        out[label] = {"rc": p.returncode, "wall_s_whole_process": round(time.time() - t0, 2), "query_phase_s": float(m.group(1)) if m else None,
                      "reads_per_s_query_phase": float(m.group(2)) if m else None, "index_load_s": float(ld.group(1)) if ld else None}
    print(json.dumps(out))
    import shutil; shutil.rmtree(tmp, ignore_errors=True)
if __name__ == "__main__":
    main()
