#!/usr/bin/env python3
"""Block-boundary statistics of the bench index: how many run-block starts fall into one directory bucket
(position-weighted, i.e. the chance that a uniformly drawn BWT position sits in such a bucket)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench, torch
from __graft_entry__ import load_package
def main():
    args = bench.parse_args()
    load_package()
    from pangenome_index_b200 import synth
    d = synth.build(args.ref_len, args.n_hap, seed=3, device=0, log=bench.log)
    ln = d["rl_len"]
    ln = (ln if torch.is_tensor(ln) else torch.as_tensor(ln)).cuda()
    ln = ln.view(torch.int64) if ln.dtype == torch.uint64 else ln.to(torch.int64)
    n = int(ln.sum()); out = {"n": n, "runs": int(ln.numel())}
    hist = torch.bincount(torch.clamp(ln, max=64), minlength=65).cpu().tolist()
    out["run_len_hist_1_8"] = hist[1:9]; out["run_len_ge_64"] = hist[64]
    starts_all = torch.cumsum(ln, 0) - ln
    for rpb in (12, 15):
        st = starts_all[::rpb]
        for sh in (6, 7, 8):
            cnt = torch.bincount(st >> sh, minlength=(n >> sh) + 1)
            # boundaries strictly inside the bucket = starts not at offset 0
            inside = torch.bincount((st[(st & ((1 << sh) - 1)) != 0]) >> sh, minlength=(n >> sh) + 1)
            h = torch.bincount(torch.clamp(inside, max=12), minlength=13).double(); h /= h.sum()
            out["rpb%d_shift%d" % (rpb, sh)] = {"frac_gt2": float(h[3:].sum()), "frac_gt4": float(h[5:].sum()), "frac_gt6": float(h[7:].sum()), "hist": [round(float(x), 5) for x in h]}
    print(json.dumps(out))
if __name__ == "__main__":
    main()
