#!/usr/bin/env python3
"""Does fmd_mem_kernel's time depend on WHERE the driver placed the index? One process, one GPU: the index is created,
timed on the bench batches, freed and created again several times; between two rounds a ballast allocation of varying size
is held so that the allocator hands out different physical pages. Prints K1 ms (timed alone by in-stream events) per round.
   python tools/placement_probe.py [bench.py flags]"""
import json, os, sys
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
    L = args.read_len; n = args.reads_per_step; pool = w["reads"].shape[0]
    d_off = torch.arange(n + 1, dtype=torch.int64, device="cuda") * L
    d_reads = [torch.cat([w["reads"][i].reshape(-1), torch.zeros(16, dtype=torch.uint8, device="cuda")]) for i in range(pool)]
    del w["reads"]; torch.cuda.empty_cache()
    out = []
    sampler = None
    if os.environ.get("PGI_PROBE_SAMPLER"):                 # does polling NVML during the run change the kernel's time?
        sampler = bench.ClockSampler(0, 0); sampler.start()
    rounds = [0, 0, 1, 0] if sampler else [0, 0, 1, 3, 7, 0, 20, 0]
    for rnd, ballast_gb in enumerate(rounds):
        ballast = torch.empty(int(ballast_gb * (1 << 30)), dtype=torch.uint8, device="cuda") if ballast_gb else None
        ix = capi.Index(w["ri"], w["tags"], w["layout"], device=0)
        batch = capi.Batch(ix, stream=torch.cuda.current_stream().cuda_stream)
        ms = []
        for i in range(3 + 6):
            batch.set_device_reads(d_reads[i % pool].data_ptr(), d_off.data_ptr(), n, n * L)
            batch.run(args.min_len, args.min_occ)
            if i >= 3: ms.append(batch.stage_ms()[0])
        out.append({"round": rnd, "ballast_gb": ballast_gb, "k1_ms_median": round(float(np.median(ms)), 3), "k1_ms_min": round(float(min(ms)), 3)})
        print("[placement]", out[-1], file=sys.stderr, flush=True)
        batch.close(); ix.close(); del ballast
        torch.cuda.empty_cache()
    if sampler: sampler.stop_flag = True
    print(json.dumps({"reads_per_step": n, "rounds": out, "clocks": sampler.summary() if sampler else None}))

if __name__ == "__main__":
    main()
