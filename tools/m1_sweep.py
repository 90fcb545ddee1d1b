#!/usr/bin/env python3
"""M1: random-access roofline sweep (independent and dependent-chain gathers) over footprints and access sizes."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from __graft_entry__ import load_package
capi = load_package().capi
out = {}
for mb in (64, 512, 2048, 16384):
    for b in (32, 64, 128):
        for chain in (1, 8):
            g = capi.random_access_gbs(0, mb << 20, b, 1 << 28, chain)
            out["%dMB_%dB_chain%d" % (mb, b, chain)] = round(g, 1)
            print("footprint %6d MB  access %3d B  chain %d : %8.1f GB/s  (%.2f G accesses/s)" % (mb, b, chain, g, g / b), flush=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "m1_sweep.json"), "w"), indent=1)
