#!/usr/bin/env python3
"""bench.py — find_mems reads/s (150 bp, MEM finding + tag lookup) on N B200s, one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the hot path (fmd_mem_kernel + tag kernels) over one batch of synthetic reads on every GPU
(weak scaling: the per-GPU batch is fixed; the index is replicated). Prints ONE JSON line on rank 0.
`value` = reads/s with the batch resident in HBM; `e2e` = the same through the C ABI with pinned HOST buffers
(H2D of reads and D2H of every result array inside the timed region). `--impl reference` times the CPU oracle
(restatement of the reference's find_mems, OpenMP over reads) on the host cores; only that leg and the
`cpu_baseline` / parity-sample legs execute anything under oracle/.
"""
import argparse, json, os, subprocess, sys, threading, time
import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

# stdout carries exactly ONE line, the JSON: file descriptor 1 is pointed at stderr for the whole run (NCCL, for one, prints
# its version banner to stdout when the first communicator comes up) and the line is written to the saved descriptor.
_JSON_OUT = None
def emit_json(line):
    out = _JSON_OUT if _JSON_OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n"); out.flush()
def capture_stdout():
    global _JSON_OUT
    if _JSON_OUT is None:
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w"); os.dup2(2, 1)

def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)

WORKLOADS = {
    # name: (ref_len, n_hap, read_len, reads_per_step, substitution rate, 1-bp indel rate, BASELINE.json config it is)
    "config4": (150_000_000, 256, 150, 2_000_000, 0.01, 0.0, "configs[3]: synthetic 256-haplotype 150 Mbp pangenome + reverse complements (random reference with N stretches, SNP 1/1000, indel 1/10000; exact BWT from the graph structure), 150 bp reads with 1 % substitutions"),
    "config3": (50_000_000, 64, 150, 2_000_000, 0.01, 0.0, "configs[2]: synthetic 64-haplotype 50 Mbp pangenome + reverse complements (random reference with N stretches, SNP 1/1000, indel 1/10000; exact BWT from the graph structure), 150 bp reads with 1 % substitutions"),
    "long": (150_000_000, 256, 10_000, 160_000, 0.03, 0.02, "configs[4]: long-read stress, 10 kbp reads with 5 % errors (3 % substitutions + 2 % 1-bp indels) on the 256-haplotype 150 Mbp index"),
}

def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=9)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--workload", default=os.environ.get("PGI_BENCH_WORKLOAD", "config4"), choices=["config4", "config3", "long", "xy", "x"])
    p.add_argument("--ref-len", type=int, default=int(os.environ.get("PGI_BENCH_REF_LEN", 0)), help="override the workload's reference length (diagnostic runs)")
    p.add_argument("--n-hap", type=int, default=int(os.environ.get("PGI_BENCH_N_HAP", 0)))
    p.add_argument("--reads-per-step", type=int, default=int(os.environ.get("PGI_BENCH_READS", 0)))
    p.add_argument("--min-len", type=int, default=20)
    p.add_argument("--min-occ", type=int, default=1)
    p.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    p.add_argument("--no-extras", action="store_true", help="skip the next-row / text-format / microbenchmark legs (scaling runs)")
    p.add_argument("--device-only", action="store_true", help="stop after the device-resident steps (profiler runs: ncu replays every kernel)")
    a = p.parse_args()
    if a.workload in WORKLOADS:
        w = WORKLOADS[a.workload]
        a.nominal = not (a.ref_len or a.n_hap)
        a.ref_len = a.ref_len or w[0]; a.n_hap = a.n_hap or w[1]; a.read_len = w[2]; a.reads_per_step = a.reads_per_step or w[3]; a.sub = w[4]; a.indel = w[5]
    else:
        a.nominal = True; a.read_len = 150; a.reads_per_step = a.reads_per_step or (1_000_000 if a.workload == "xy" else 10_000); a.sub = 0.01 if a.workload == "xy" else 0.0; a.indel = 0.0
    return a

class ClockSampler(threading.Thread):
    """SM clocks / throttle reasons during the timed region (B200_PROFILING.md), read in-process through NVML
    (nvidia_ml_py). Spawning `nvidia-smi` every 0.2 s from each of 8 ranks enumerated all GPUs under the driver's global
    lock and stretched the host side of every step (r01o: 19.6 ms per step on 8 GPUs against 12.3 ms on one);
    nvidia-smi remains the fallback, on rank 0 only, at a 1 s period."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    def __init__(self, gpu, rank=0):
        super().__init__(daemon=True); self.gpu = gpu; self.rank = rank; self.rows = []; self.stop_flag = False; self.h = None; self.nv = None; self.extra = []
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            idx = gpu
            if vis and all(x.strip().isdigit() for x in vis.split(",")): idx = int(vis.split(",")[gpu])
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx); self.nv = pynvml
        except Exception:
            self.h = None
    def sample_nvml(self):
        nv = self.nv
        sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM); mx = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
        get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        r = int(get(self.h))
        bit = lambda *names: next((getattr(nv, n) for n in names if hasattr(nv, n)), 0)
        flags = [bit("nvmlClocksEventReasonHwSlowdown", "nvmlClocksThrottleReasonHwSlowdown"),
                 bit("nvmlClocksEventReasonHwThermalSlowdown", "nvmlClocksThrottleReasonHwThermalSlowdown"),
                 bit("nvmlClocksEventReasonSwThermalSlowdown", "nvmlClocksThrottleReasonSwThermalSlowdown"),
                 bit("nvmlClocksEventReasonSwPowerCap", "nvmlClocksThrottleReasonSwPowerCap")]
        try:
            temp = nv.nvmlDeviceGetTemperature(self.h, nv.NVML_TEMPERATURE_GPU); mtemp = None
            try:
                fv = nv.nvmlDeviceGetFieldValues(self.h, [nv.NVML_FI_DEV_MEMORY_TEMP])[0]
                if fv.nvmlReturn == 0: mtemp = int(fv.value.uiVal)
            except Exception:
                pass
            self.extra.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_MEM), nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0, r, temp, mtemp))
        except Exception:
            pass
        return [str(sm), str(mx)] + ["Active" if (r & f) else "Not Active" for f in flags]
    def run(self):
        while not self.stop_flag:
            try:
                if self.h is not None:
                    self.rows.append(self.sample_nvml()); time.sleep(0.1); continue
                if self.rank == 0:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                         capture_output=True, text=True, timeout=5).stdout.strip()
                    if out: self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(1.0)
    def summary(self):
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        out = {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(self.rows)}
        if self.extra:
            out["mem_mhz"] = float(np.median([e[0] for e in self.extra])); out["power_w_max"] = round(max(e[1] for e in self.extra), 1)
            m = 0
            for e in self.extra: m |= e[2]
            out["reasons_mask_or"] = hex(m)
            out["gpu_temp_c_max"] = max(e[3] for e in self.extra)
            mt = [e[4] for e in self.extra if e[4] is not None]
            out["hbm_temp_c_max"] = max(mt) if mt else None
        return out

E2E_CALLERS = 3

def host_threads():
    """Host threads the CPU legs use: every core this process may run on. Passed to the oracle explicitly, because
    torchrun exports OMP_NUM_THREADS=1 to its workers and omp_get_max_threads() would then say 1."""
    try: return max(1, len(os.sched_getaffinity(0)))
    except Exception: return max(1, os.cpu_count() or 1)

class Workload:
    """Index + read batches for one rank. Data prep only (never timed)."""
    pass

def files_for_oracle(w, prep):
    """.ri / tags bytes of the workload, for the CPU legs (the reference's own loaders read these)."""
    if getattr(w, "ri", None) is None:
        t0 = time.time()
        h = w.host_runs()
        w.ri, info = prep.rindex_from_rlbwt(prep.rlbwt_from_runs(h["rl_sym"], h["rl_len"]), encoded=True, with_samples=False)
        w.tags, _ = prep.tags_from_runs(h["tag_val"], h["tag_len"], 0); w.layout = 0
        w.facts.update(ri_bytes=len(w.ri), tags_bytes=len(w.tags), files_for_cpu_legs_s=round(time.time() - t0, 1))
    return w.ri, w.tags, w.layout

def make_workload(args, dev_index, rank, world=1, dist=None, need_index=True):
    """-> Workload(ix or None, reads (cuda uint8 [pool, n, L]), name, facts, host_runs()). Multi-GPU: rank 0 builds the
    run-length BWT + tag runs ONCE per node and broadcasts them over NCCL (NVLink); every rank builds its own replica of the
    index from the runs on its own GPU (pgi_index_create_from_runs)."""
    import torch
    from __graft_entry__ import load_package
    pkg = load_package(); prep = pkg.prep; capi = pkg.capi
    dev = torch.device("cuda", dev_index)
    w = Workload(); t0 = time.time()
    pool = max(1, min(4, args.steps + args.warmup)); n = args.reads_per_step; L = args.read_len
    if args.workload in ("xy", "x"):
        import refcases, datasets
        from pangenome_index_b200 import synth
        if args.workload == "xy":
            w.ri = datasets.gold("xy.ri"); w.tags = datasets.gold("xy_bidirectional_compressed.tags"); w.layout = 1; text = datasets.gold("contigs_xy")
            w.name = "configs[1]: shipped bidirectional_test xy.ri (legacy, n=8022, sigma=5) + xy_bidirectional_compressed.tags (ByteCode), reads with 1 % substitutions"
        else:
            x = refcases.x_index(); w.ri, w.tags, w.layout, text = x["ri"], x["tags"], 0, x["text"]
            w.name = "configs[0]: index rebuilt from the shipped x.rl_bwt (encoded, n=3012, sigma=5, forward-only) + synthetic aligned tags, error-free reads"
        tt = torch.frombuffer(bytearray(text), dtype=torch.uint8).to(dev)
        starts = np.concatenate([[0], np.nonzero(np.frombuffer(text, dtype=np.uint8) == 10)[0] + 1])
        w.reads = torch.stack([synth.sample_reads(tt, starts, n, L, args.sub, 2 + 97 * rank + 7 * i, dev) for i in range(pool)])
        w.facts = {"n": len(text)}
        w.ix = capi.Index(w.ri, w.tags, w.layout, device=dev_index) if need_index else None
        w.index_create_s = 0.0
        return w
    from pangenome_index_b200 import pangraph as pg
    lg = log if rank == 0 else None
    prepared = pg.prepare(args.ref_len, args.n_hap, seed=3, device=dev_index, log=lg)      # every rank: the graph (reads are sampled from it)
    g = prepared[0]
    t1 = time.time()
    runs = None
    if rank == 0:
        d = pg.build(args.ref_len, args.n_hap, prepared=prepared, log=lg, host=False); runs = d["dev_runs"]; del d
    bcast_s = 0.0; bcast_bytes = 0
    if world > 1:
        torch.cuda.synchronize(); dist.barrier(); tb = time.time()
        runs, bcast_bytes = pkg.sharding.broadcast_runs(runs, dev)
        torch.cuda.synchronize(); bcast_s = time.time() - tb
    build_s = time.time() - t1
    w.host_runs = lambda: pg.host_runs(runs)
    n_bwt = int(runs["len"].sum())
    t2 = time.time()
    w.ix = capi.Index.from_runs(runs["sym"], runs["len"], runs["tag_val"], runs["tag_len"], device=dev_index) if need_index else None
    torch.cuda.synchronize(); w.index_create_s = time.time() - t2
    w.reads = torch.stack([pg.sample_reads(prepared[1], n, L, args.sub, 4 + 97 * rank + 7 * i, indel=args.indel) for i in range(pool)])
    w.runs = runs
    w.facts = {"n": n_bwt, "r": int(len(runs["sym"])), "sequences": 2 * g.H, "haplotypes": g.H, "reference_bp": g.L, "sites": g.S, "tag_runs": int(len(runs["tag_val"])),
               "graph_and_bwt_s": round(build_s - bcast_s, 1), "nccl_broadcast_of_runs_s": round(bcast_s, 3), "runs_broadcast_bytes": int(bcast_bytes),
               "index_from_runs_on_gpu_s": round(w.index_create_s, 2), "prep_s": round(time.time() - t0, 1)}
    w.name = WORKLOADS[args.workload][6] if args.nominal else ("%s at a NON-NOMINAL size: %d haplotypes x %.1f Mbp" % (args.workload, args.n_hap, args.ref_len / 1e6))
    w.ri = None
    if rank == 0: log("workload ready:", w.facts)
    return w

def cpu_rate(o, R, reads_np, L, min_len, min_occ, target_s, threads=0):
    """Time the CPU baseline on a bounded sample sized for about target_s seconds: the reference's own code (R, oracle/_ref)
    when it is built, else the restatement (o). -> dict(reads, seconds, rate, kind)"""
    threads = threads or host_threads()
    n_avail = len(reads_np) // L
    def run(n):
        off = np.arange(n + 1, dtype=np.uint64) * np.uint64(L)
        if R is not None: return R.ref_time(reads_np[: n * L], off, min_len, min_occ, threads=threads)[0]
        return o.find_mems(reads_np[: n * L], off, min_len, min_occ, threads=threads, text=False)["seconds"]
    probe = min(n_avail, max(64, 3_000_000 // L))
    dt = max(1e-6, run(probe))
    n = int(min(n_avail, max(probe, probe * target_s / dt)))
    dt = run(n)
    return dict(reads=n, seconds=dt, rate=n / dt, kind="reference" if R is not None else "port")

def cpu_indexes(w, prep):
    """(oracle restatement, reference library or None) over the workload's files."""
    import orc
    ri, tags, layout = files_for_oracle(w, prep)
    o = orc.Index(ri, tags, layout)
    R = None
    try:
        if orc.ref_available(): R = orc.Index(ri, tags, layout, impl="ref")
    except Exception as e:
        log("reference library not usable, falling back to the restatement:", e)
    return o, R

def run_reference(args, rank, world):
    """--impl reference: the reference's own find_mems code on the host cores (oracle/_ref/libpgiref.so = src/r-index.cpp +
    src/tag_arrays.cpp + find_all_mems compiled against stand-in SDSL headers; OpenMP over reads is ours -- the reference's
    loop is serial), or the restatement in oracle/ when that library is not built. Rank 0 only."""
    if rank != 0:
        return
    import torch
    from __graft_entry__ import load_package
    prep = load_package().prep
    w = make_workload(args, 0, 0, need_index=False)
    o, R = cpu_indexes(w, prep)
    L = args.read_len
    reads_np = w.reads[0].reshape(-1).cpu().numpy()
    cores = host_threads()
    per_step_s = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    probe = cpu_rate(o, R, reads_np, L, args.min_len, args.min_occ, per_step_s)
    n = probe["reads"]; off = np.arange(n + 1, dtype=np.uint64) * np.uint64(L)
    def step():
        if R is not None: return R.ref_time(reads_np[: n * L], off, args.min_len, args.min_occ, threads=cores)[0]
        return o.find_mems(reads_np[: n * L], off, args.min_len, args.min_occ, threads=cores, text=False)["seconds"]
    for _ in range(max(0, args.warmup - 1)): step()
    t = sum(step() for _ in range(args.steps))
    rate = n * args.steps / t
    kind = probe["kind"]
    sample = "%d reads per step (bounded sample of the %d-read batch), %d OpenMP threads" % (n, args.reads_per_step, cores)
    note = ("the reference's own code: src/r-index.cpp + src/tag_arrays.cpp compiled unmodified against stand-in SDSL / GBWT headers, find_all_mems + tag query per MEM "
            "cut out by line range (oracle/Makefile); its find_mems loop is single-threaded, the OpenMP loop over reads is ours; stdout muted") if kind == "reference" else \
           "oracle/_ref is not built on this machine: the CPU restatement in oracle/, parallelised over reads with OpenMP"
    line = {"impl": "reference", "metric": "find_mems reads/s (150bp, MEM + tag query)", "value": rate, "unit": "reads/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": {"workload": w.name, "workload_key": args.workload, "facts": w.facts, "reads_per_step_per_gpu": args.reads_per_step, "read_len": L, "substitution_rate": args.sub, "indel_rate": args.indel,
                       "min_len": args.min_len, "min_occ": args.min_occ, "sample_reads_per_step": n, "note": note},
            "cpu_baseline": {"value": rate, "unit": "reads/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": rate, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit_json(line)

def pin_rank_to_own_cores(local, world):
    """One process per GPU on DISJOINT cores: the cores NVML names for this GPU (its NUMA node) are divided among the ranks
    that share them, so that a rank's stream synchronisation never waits behind another rank's threads (SCALE_r01: all ranks
    ran on cores 0-31 and one of them lost 25 % to a single 25 ms host hiccup). Returns the original affinity mask."""
    orig = os.sched_getaffinity(0)
    if world <= 1: return orig
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", ""); idx = local
        if vis and all(x.strip().isdigit() for x in vis.split(",")): idx = int(vis.split(",")[local])
        words = pynvml.nvmlDeviceGetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(idx), (os.cpu_count() + 63) // 64)
        near = sorted(c for c in orig if (words[c // 64] >> (c % 64)) & 1) or sorted(orig)
    except Exception as e:
        log("NVML affinity unavailable:", e); near = sorted(orig)
    # ranks sharing the same `near` set split it evenly (ranks are mapped to GPUs in order: neighbours share a node)
    sharers = max(1, min(world, round(world * len(near) / max(1, len(orig)))))
    per = max(1, len(near) // sharers); k = local % sharers
    mine = near[k * per:(k + 1) * per] or near
    try: os.sched_setaffinity(0, mine)
    except Exception as e: log("cpu affinity not set:", e)
    return orig

def main():
    capture_stdout()
    args = parse_args()
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank, world)
    import torch, torch.distributed as dist
    from __graft_entry__ import load_package
    pkg = load_package(); capi = pkg.capi; prep = pkg.prep
    if capi.lib().pgi_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the find_mems path has no CPU fallback")
    torch.cuda.set_device(local)
    orig_affinity = pin_rank_to_own_cores(local, world)
    my_cores = sorted(os.sched_getaffinity(0))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    def barrier():
        if world > 1: dist.barrier()
        torch.cuda.synchronize()
    l2_fetch = capi.l2_fetch_granularity(local, int(os.environ.get("PGI_L2FETCH", 0)))
    w = make_workload(args, local, rank, world, dist)
    ix = w.ix
    if rank == 0: log("index on device: %.1f MB in %.2fs" % (ix.info["device_bytes"] / 1e6, w.index_create_s), ix.info)
    L = args.read_len; n = args.reads_per_step; pool = w.reads.shape[0]
    d_off = torch.arange(n + 1, dtype=torch.int64, device="cuda") * L
    stream = torch.cuda.current_stream()
    batch = capi.Batch(ix, stream=stream.cuda_stream)
    # pinned host copies of the read batches for the end-to-end leg
    h_reads = [torch.empty(n * L + 16, dtype=torch.uint8).pin_memory() for _ in range(pool)]
    for i in range(pool): h_reads[i][: n * L].copy_(w.reads[i].reshape(-1))
    h_off = (torch.arange(n + 1, dtype=torch.int64) * L).pin_memory()
    d_reads = [torch.cat([w.reads[i].reshape(-1), torch.zeros(16, dtype=torch.uint8, device="cuda")]) for i in range(pool)]
    del w.reads
    torch.cuda.synchronize()

    def step_device(i):
        batch.set_device_reads(d_reads[i % pool].data_ptr(), d_off.data_ptr(), n, n * L)
        batch.run(args.min_len, args.min_occ)
    def step_e2e(i):
        """The public call: pgi_find_mems with pinned HOST buffers in, every result array in pinned HOST memory out."""
        r = ix.find_mems_raw(h_reads[i % pool].data_ptr(), h_off.data_ptr(), n, args.min_len, args.min_occ)
        sizes = (r.n_reads, r.n_mems, r.n_tags)
        ix.result_free(r)
        return sizes

    # ---- device-resident timing (value)
    for i in range(args.warmup): step_device(i)
    sampler = ClockSampler(local, rank)
    if not os.environ.get("PGI_BENCH_NOSAMPLER"): sampler.start()
    barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    k1_ms = 0.0; stage = np.zeros(8); ext_total = 0; launches = 0; mems_total = 0; tags_total = 0; seed_total = 0; blk_total = 0; reruns = 0
    e0.record(stream); host_gap = []; t_prev = time.perf_counter()
    for i in range(args.steps):
        step_device(args.warmup + i)
        t_now = time.perf_counter(); host_gap.append(1e3 * (t_now - t_prev)); t_prev = t_now      # host wall time of the step (diagnostic)
        ms = batch.stage_ms(); stage += np.array(ms); k1_ms += ms[0]; launches += int(ms[6]); reruns += int(ms[7])
        r = batch.device_result(); ext_total += r.extensions; mems_total += r.n_mems; tags_total += r.n_tags; seed_total += r.seed_lookups; blk_total += r.blocks_fetched
    e1.record(stream)
    barrier()
    dev_ms = e0.elapsed_time(e1)
    if args.device_only:
        log("device-only run: %.3f ms per step, fmd_mem_kernel %.3f ms, L2 fetch granularity %d B, K1 variant %s" % (dev_ms / args.steps, k1_ms / args.steps, l2_fetch, os.environ.get("PGI_K1", "default")))
        sampler.stop_flag = True
        return
    # ---- the same steps with TWO batches in flight from this one host thread (pgi_batch_run_async / pgi_batch_wait, only the
    # fmd_mem_kernel launches ordered on the GPU: PGI_CHAIN=3): the next step's persistent grid takes the SMs that the previous
    # one's last reads leave idle and the short ordering / tag kernels run beside it. Reported next to `value`, not as it: the
    # per-kernel timings of the roofline object need one step at a time.
    pipelined = None
    if not args.no_extras:
        old_chain = os.environ.get("PGI_CHAIN"); os.environ["PGI_CHAIN"] = "3"
        b2 = [batch, capi.Batch(ix)]
        def run_pipe(k_steps, first):
            fl = []
            for i in range(k_steps):
                b = b2[i % 2]
                if len(fl) == 2: fl.pop(0).wait()
                b.set_device_reads(d_reads[(first + i) % pool].data_ptr(), d_off.data_ptr(), n, n * L); b.run_async(args.min_len, args.min_occ); fl.append(b)
            for b in fl: b.wait()
        run_pipe(max(2, args.warmup), 0); barrier()
        tp = time.perf_counter(); run_pipe(args.steps, args.warmup); torch.cuda.synchronize(); pipe_ms = 1e3 * (time.perf_counter() - tp)
        barrier()
        b2[1].close()
        if old_chain is None: os.environ.pop("PGI_CHAIN", None)
        else: os.environ["PGI_CHAIN"] = old_chain
        pipelined = {"value": n * args.steps / (pipe_ms * 1e-3), "unit": "reads/s per GPU (rank 0)", "ms_per_step": pipe_ms / args.steps, "batches_in_flight": 2,
                     "note": "pgi_batch_run_async / pgi_batch_wait, fmd_mem_kernel launches chained on the GPU (PGI_CHAIN=3); host wall clock"}
    # ---- end-to-end timing: K calls of pgi_find_mems (pinned host buffers in, all result arrays out to pinned host memory),
    # issued by E2E_CALLERS host threads the way the batched find_mems driver does, so the copies of one call overlap the
    # kernels of another. Wall clock around all K calls (every call blocks until its result is in host memory).
    def step_e2e_compact(i):
        """The lean form of the same call: fixed-length reads (no offset array), result packed on the GPU (pgi_find_mems_compact)."""
        r = ix.find_mems_compact_raw(h_reads[i % pool].data_ptr(), None, L, n, args.min_len, args.min_occ)
        sizes = (r.n_reads, r.n_mems, r.n_tags, r.tag_bytes)
        ix.result_compact_free(r)
        return sizes
    def e2e_pass(k_calls, first, compact):
        nxt = [first]; lock = threading.Lock(); acc = [0]
        def caller():
            while True:
                with lock:
                    i = nxt[0]
                    if i >= first + k_calls: return
                    nxt[0] += 1
                if compact:
                    nr_, nm_, nt_, tb_ = step_e2e_compact(i); got = nr_ * 4 + nm_ * 24 + nt_ * tb_
                else:
                    nr_, nm_, nt_ = step_e2e(i); got = (nr_ + 1) * 8 + nm_ * 32 + (nm_ + 1) * 8 + nt_ * 8 + nm_ * 8
                with lock: acc[0] += got
        ths = [threading.Thread(target=caller) for _ in range(min(E2E_CALLERS, k_calls))]
        t0 = time.perf_counter(); [t.start() for t in ths]; [t.join() for t in ths]
        return 1e3 * (time.perf_counter() - t0), acc[0]
    e2e_pass(max(args.warmup, E2E_CALLERS), 0, False)
    barrier()
    e2e_full_ms, d2h_full = e2e_pass(args.steps, args.warmup, False)
    barrier()
    e2e_pass(max(args.warmup, E2E_CALLERS), 0, True)
    barrier()
    e2e_ms, d2h = e2e_pass(args.steps, args.warmup, True)
    barrier()
    sampler.stop_flag = True
    if sampler.is_alive(): sampler.join(timeout=2)
    # ---- max over ranks
    t = torch.tensor([dev_ms, e2e_ms, e2e_full_ms], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([launches, ext_total, mems_total, tags_total], dtype=torch.float64, device="cuda")
    per_rank = [[dev_ms / args.steps, e2e_ms / args.steps, max(host_gap), k1_ms / args.steps, e2e_full_ms / args.steps]]
    if world > 1:
        mine = torch.tensor(per_rank[0], dtype=torch.float64, device="cuda"); allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine); per_rank = [[round(float(x), 3) for x in r] for r in allr]
        dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    dev_ms, e2e_ms, e2e_full_ms = float(t[0]), float(t[1]), float(t[2])
    if rank == 0: log("per rank [device ms/step, e2e ms/step, slowest host step ms, K1 ms]:", per_rank, "rank 0 host ms per step:", [round(x, 2) for x in host_gap], "cores of rank 0:", my_cores)
    if rank != 0:
        if world > 1: dist.destroy_process_group()
        return
    total_reads = n * args.steps * world
    value = total_reads / (dev_ms * 1e-3); e2e = total_reads / (e2e_ms * 1e-3); e2e_full = total_reads / (e2e_full_ms * 1e-3)
    try: os.sched_setaffinity(0, orig_affinity)          # the CPU legs use every host core again
    except Exception: pass
    # ---- CPU side (rank 0, never inside a timed region): the files the reference's loaders read, the restatement, the reference library
    o, R = cpu_indexes(w, prep)
    reads_np = h_reads[0][: n * L].numpy()
    # parity of a FULL timed batch: the same batch object and the same 2 M-read device batch as the timed steps (partitioned
    # MEM slots and all); its first `ns` reads are compared with the oracle array by array, the whole result is summarised.
    ns = min(n, max(2000, min(200_000, 30_000_000 // L))); offs = np.arange(ns + 1, dtype=np.uint64) * np.uint64(L)
    batch.set_device_reads(d_reads[0].data_ptr(), d_off.data_ptr(), n, n * L); batch.run(args.min_len, args.min_occ)
    G = batch.fetch(text=False)
    Ro = o.find_mems(reads_np[: ns * L], offs, args.min_len, args.min_occ, count_work=True, text=False)
    m1 = int(G["mem_off"][ns]); t1 = int(G["tag_off"][m1])
    parity = bool(m1 == Ro["n_mems"] and np.array_equal(G["mem_off"][: ns + 1], Ro["mem_off"]) and np.array_equal(G["mems"][:m1], Ro["mems"])
                  and np.array_equal(G["tag_off"][: m1 + 1], Ro["tag_off"]) and np.array_equal(G["tags"][:t1], Ro["tags"]) and np.array_equal(G["n_runs"][:m1], Ro["n_runs"]))
    parity_ref = None
    if R is not None:                                       # and a slice against the reference's own code
        nr = min(ns, 20000); Rr = R.find_mems(reads_np[: nr * L], offs[: nr + 1], args.min_len, args.min_occ, text=False, threads=host_threads())
        mr = int(G["mem_off"][nr]); tr = int(G["tag_off"][mr])
        parity_ref = bool(np.array_equal(G["mems"][:mr], Rr["mems"]) and np.array_equal(G["tags"][:tr], Rr["tags"]) and np.array_equal(G["n_runs"][:mr], Rr["n_runs"]))
    small = ix.find_mems(reads_np[: 2000 * L] if n >= 2000 else reads_np, offs[: min(n, 2000) + 1], args.min_len, args.min_occ, text=False)   # small batches take the single-counter path
    parity_small = bool(np.array_equal(small["mems"], G["mems"][: small["n_mems"]]))
    wide = {"max_bwt_start": int(G["mems"][:, 2].max()) if G["n_mems"] else 0, "mems_beyond_2^32": int((G["mems"][:, 2] >= (1 << 32)).sum())}
    import hashlib
    batch_digest = hashlib.sha256(G["mems"].tobytes() + G["tags"].tobytes() + G["n_runs"].tobytes()).hexdigest()[:16]
    # ---- next rows of the scope table (SURVEY.md section 8 f), reported beside the headline, not part of it
    next_rows = {}; fmt = {}; rnd = {}
    if not args.no_extras:
        try:
            qs = []
            for i in range(2 + 3):
                batch.set_device_reads(d_reads[i % pool].data_ptr(), d_off.data_ptr(), n, n * L)
                batch.run_query_tags()
                if i >= 2: qs.append((batch.stage_ms()[5], batch.device_result().extensions))
            nq = min(ns, 20000)
            QG = ix.query_tags(reads_np[: nq * L], offs[: nq + 1], text=False); QR = o.query_tags(reads_np[: nq * L], offs[: nq + 1], text=False, threads=host_threads())
            q_par = bool(np.array_equal(QG["mems"], QR["mems"]) and np.array_equal(QG["tags"], QR["tags"]) and np.array_equal(QG["n_runs"], QR["n_runs"]))
            q_ms = float(np.median([x[0] for x in qs]))
            next_rows["query_tags"] = {"reads_per_s": n / (q_ms * 1e-3), "ms_per_step": q_ms, "lf_steps_per_read": qs[-1][1] / n,
                                       "matched_fraction": float((QR["mems"][:, 3] > 0).mean()), "parity_sample": "%d reads bit-exact vs oracle: %s" % (nq, q_par),
                                       "cpu_port_reads_per_s": nq / QR["seconds"], "note": "reads searched whole (right to left until the range is empty)"}
        except Exception as e:
            next_rows["query_tags"] = {"error": str(e)}
        # text formatting (src/find_mems.cpp:115-139 stdout grammar): host-bound, reported separately (SURVEY section 8 d)
        try:
            import ctypes
            r = ix.find_mems_raw(h_reads[0].data_ptr(), h_off.data_ptr(), n, args.min_len, args.min_occ)
            ln = ctypes.c_uint64(); t1_ = time.perf_counter()
            ptxt = capi.lib().pgi_format_text(ctypes.byref(r), 1, 1, ctypes.byref(ln)); t_fmt = time.perf_counter() - t1_
            capi.lib().pgi_free(ptxt); ix.result_free(r)
            fmt = {"reads_per_s": n / t_fmt, "text_bytes_per_step": int(ln.value), "threads": host_threads(), "note": "pgi_format_text (read ranges formatted in parallel); not inside value / e2e"}
        except Exception as e:
            fmt = {"error": str(e)}
        # M1: random-access figures over a buffer of the size of the run-block array (capped), for the record
        try:
            foot = int(min(max(ix.info["n_dev_blocks"] * 64, 1 << 30), 24 << 30))
            for b in (32, 64):
                rnd["random%d_ceiling_gbs" % b] = round(capi.random_access_gbs(local, foot, b, 1 << 28, 0), 1)
            rnd["random64_chained8_gbs"] = round(capi.random_access_gbs(local, foot, 64, 1 << 26, 8), 1)
            rnd["footprint_bytes"] = foot
        except Exception as e:
            rnd["error"] = str(e)
    c = Ro["counters"]; d_per_e = c["D"] / max(1, c["E"])
    cpu = cpu_rate(o, R, reads_np, L, args.min_len, args.min_occ, args.cpu_seconds)
    cores = host_threads()
    # ---- roofline of the dominant kernel (fmd_mem_kernel)
    peaks = {}
    try: peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception: pass
    peak = float(peaks.get("hbm_gbs", 6650.0)); peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    k1_s = (k1_ms / args.steps) * 1e-3
    # (a) what the kernel MOVES, counted by the kernel itself: 64 B per run block fetched + 16 B per seed lookup + the read
    # bytes and their 2-bit codes + 48 B per MEM appended (32 B record, 8 B key, 8 B per-read count share)
    blk = blk_total / args.steps; seeds = seed_total / args.steps; mems_step = mems_total / args.steps
    moved = 64.0 * blk + 16.0 * seeds + n * L * 1.25 + 48.0 * mems_step
    achieved = moved / k1_s / 1e9
    # (b) the REFERENCE algorithm's bytes for the same answers (SURVEY.md section 8 d): 128 B per distinct 10-run block per extension
    equiv = 128.0 * d_per_e * (ext_total / args.steps)
    traffic = None; traffic_src = None
    try:                                                    # ncu dram bytes per launch, only from a capture of THIS kernel build on THIS workload
        import hashlib as _h
        hh = _h.sha256()
        for f in sorted(os.listdir(os.path.join(ROOT, "pangenome-index_b200", "csrc"))):
            if f in ("pgi_kernels.cuh", "pgi_core.cuh", "pgi_layout.h"): hh.update(open(os.path.join(ROOT, "pangenome-index_b200", "csrc", f), "rb").read())
        tj = json.load(open(os.path.join(ROOT, "profiles", "k1_traffic.json")))
        for e in tj:
            if e["csrc_sha256"] == hh.hexdigest() and e["workload"] == args.workload and e["nominal"] == args.nominal and e["reads_per_step"] == n:
                traffic = e["dram_bytes_per_launch"]; traffic_src = e["source"]
    except Exception:
        pass
    line = {"metric": "find_mems reads/s (150bp, MEM + tag query)", "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int64", "data": "synthetic",
            "config": {"workload": w.name, "workload_key": args.workload, "facts": w.facts, "reads_per_step_per_gpu": n, "read_len": L, "substitution_rate": args.sub, "indel_rate": args.indel,
                       "min_len": args.min_len, "min_occ": args.min_occ, "index": "built once per node on rank 0's GPU, run arrays broadcast over NCCL, replica constructed on every GPU from the runs",
                       "l2_fetch_granularity_bytes": l2_fetch, "device_index_bytes": ix.info["device_bytes"], "bucket_shift": ix.info["bucket_shift"], "seed_k": ix.info["seed_k"],
                       "l2_policy": "read batch (%.0f MB) and a %.1f GB index, cycled over %d distinct batches; inputs larger than L2" % (n * L / 1e6, ix.info["device_bytes"] / 1e9, pool),
                       "parity_full_batch": "first %d reads of a timed-path %d-read batch bit-exact vs oracle (MEMs, offsets, tags, n_runs): %s" % (ns, n, parity),
                       "parity_vs_reference_code": parity_ref, "parity_small_batch_path": parity_small, "batch_digest": batch_digest, "positions_beyond_32_bits": wide,
                       "mems_per_read": float(cnt[2]) / total_reads, "tags_per_read": float(cnt[3]) / total_reads,
                       "extensions_per_read": float(cnt[1]) / total_reads, "step_reruns": reruns},
            "pipelined": pipelined,
            "e2e": {"value": e2e, "unit": "reads/s", "h2d_bytes_per_step": n * L, "d2h_bytes_per_step": d2h // args.steps, "ms_per_step": e2e_ms / args.steps, "concurrent_callers": E2E_CALLERS,
                    "api": "pgi_find_mems_compact: pinned host reads in (fixed length, no offset array), every result array out to pinned host memory in the compact wire form (packed on the GPU; pgi_result_expand restores the reference-layout arrays on the host when a caller wants them)",
                    "bytes_per_read": (n * L + d2h / args.steps) / n,
                    "reference_layout": {"value": e2e_full, "unit": "reads/s", "api": "pgi_find_mems: reads + u64 offsets in, pgi_result (struct MEM records, u64 offsets / tags / n_runs) out", "h2d_bytes_per_step": n * L + (n + 1) * 8,
                                         "d2h_bytes_per_step": d2h_full // args.steps, "ms_per_step": e2e_full_ms / args.steps, "bytes_per_read": (n * L + (n + 1) * 8 + d2h_full / args.steps) / n}},
            "gpu_launches": int(cnt[0]),
            "per_rank_ms_per_step": {"device": [r[0] for r in per_rank], "e2e": [r[1] for r in per_rank], "e2e_reference_layout": [r[4] for r in per_rank], "slowest_host_step": [r[2] for r in per_rank], "fmd_mem_kernel": [r[3] for r in per_rank]},
            "stage_ms_per_step": {k: round(float(v) / args.steps, 3) for k, v in zip(["fmd_mem_kernel", "mem_order", "tag_plan", "tag_decode_sort", "tag_unique_compact", "run_total"], stage[:6])},
            "roofline": {"bound": "hbm", "kernel": "fmd_mem_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": moved,
                         "bytes_model": "bytes the kernel moves, counted by the kernel: 64 B x %d run blocks fetched + 16 B x %d seed lookups + 1.25 B x %d read bytes (+ 2-bit codes) + 48 B x %d MEMs" % (blk, seeds, n * L, mems_step),
                         "blocks_fetched_per_launch": blk, "seed_lookups_per_launch": seeds, "executed_extensions_per_launch": (ext_total - seed_total * ix.info["seed_k"]) / args.steps,
                         "reference_algorithm_equivalent": {"bytes_per_launch": equiv, "gbs": equiv / k1_s / 1e9, "over_peak": equiv / k1_s / 1e9 / peak,
                                                            "model": "SURVEY.md 8(d): 128 B x D (oracle: D/E = %.3f) x E = %d extensions of the reference algorithm per launch; NOT a bandwidth: the seed table answers %d of them from 16-byte entries and the layout moves 64 B where the reference moves 128 B" % (d_per_e, ext_total // args.steps, int(seeds * ix.info["seed_k"]))},
                         "traffic_gbs": (traffic / k1_s / 1e9) if traffic else None, "frac_traffic": (traffic / k1_s / 1e9 / peak) if traffic else None,
                         "kernel_ms_per_launch": k1_ms / args.steps, **rnd},
            "cpu_baseline": {"value": cpu["rate"], "unit": "reads/s", "cores": cores, "kind": cpu["kind"],
                             "sample": "%d reads of the same batch in %.1f s, %d OpenMP threads (%s; the reference's own loop is single-threaded)" % (cpu["reads"], cpu["seconds"], cores, "the reference's own code, oracle/_ref" if cpu["kind"] == "reference" else "oracle/ restatement")},
            "index_create_s": round(w.index_create_s, 2), "text_format": fmt,
            "next_rows": next_rows,
            "clocks": sampler.summary()}
    emit_json(line)
    if world > 1: dist.destroy_process_group()

if __name__ == "__main__":
    main()
