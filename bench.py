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

def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=9)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--workload", default="synth", choices=["synth", "xy"])
    p.add_argument("--ref-len", type=int, default=int(os.environ.get("PGI_BENCH_REF_LEN", 50_000_000)))
    p.add_argument("--n-hap", type=int, default=int(os.environ.get("PGI_BENCH_N_HAP", 20)))
    p.add_argument("--reads-per-step", type=int, default=int(os.environ.get("PGI_BENCH_READS", 2_000_000)))
    p.add_argument("--read-len", type=int, default=150)
    p.add_argument("--sub", type=float, default=0.01)
    p.add_argument("--min-len", type=int, default=20)
    p.add_argument("--min-occ", type=int, default=1)
    p.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    return p.parse_args()

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

def make_workload(args, dev_index, rank):
    """-> dict(ri, tags, layout, reads (cuda uint8 [pool, n, L]), name, facts). Data prep only (not timed)."""
    import torch
    from __graft_entry__ import load_package
    pkg = load_package(); prep = pkg.prep
    dev = torch.device("cuda", dev_index)
    t0 = time.time()
    pool = max(1, min(4, args.steps + args.warmup))
    if args.workload == "xy":
        gold = os.path.join(ROOT, "tests", "golden")
        ri = open(os.path.join(gold, "xy.ri"), "rb").read(); tags = open(os.path.join(gold, "xy_bidirectional_compressed.tags"), "rb").read()
        text = open(os.path.join(gold, "contigs_xy"), "rb").read()
        from pangenome_index_b200 import synth
        tt = torch.frombuffer(bytearray(text), dtype=torch.uint8).to(dev)
        starts = np.concatenate([[0], np.nonzero(np.frombuffer(text, dtype=np.uint8) == 10)[0] + 1])
        reads = torch.stack([synth.sample_reads(tt, starts, args.reads_per_step, args.read_len, args.sub, 2 + 97 * rank + 7 * i, dev) for i in range(pool)])
        name = "configs[1]: shipped bidirectional_test xy.ri (legacy, n=8022, sigma=5) + xy_bidirectional_compressed.tags (ByteCode)"
        return dict(ri=ri, tags=tags, layout=1, reads=reads, name=name, facts={"n": 8022})
    from pangenome_index_b200 import synth
    d = synth.build(args.ref_len, args.n_hap, seed=3, device=dev_index, log=log if rank == 0 else None)
    reads = torch.stack([synth.sample_reads(d["text"], d["seq_starts"], args.reads_per_step, args.read_len, args.sub, 4 + 97 * rank + 7 * i, dev) for i in range(pool)])
    del d["text"]; torch.cuda.empty_cache()
    rl = prep.rlbwt_from_runs(d["rl_sym"], d["rl_len"])
    ri, info = prep.rindex_from_rlbwt(rl, encoded=True, with_samples=False)
    del rl
    tags, n_tag_runs = prep.tags_from_runs(d["tag_val"], d["tag_len"], 0)
    facts = {"n": int(d["n"]), "r": int(info["total_runs"]), "sequences": int(d["n_seq"]), "tag_runs": int(n_tag_runs),
             "ri_bytes": len(ri), "tags_bytes": len(tags), "prep_s": round(time.time() - t0, 1)}
    name = ("configs[2]-style synthetic pangenome: %d haplotypes x %.1f Mbp + reverse complements (nominal config: 64 x 50 Mbp; "
            "built size stated in facts), SNP 1/1000, indel 1/10000, N stretches (sigma=6)" % (args.n_hap, args.ref_len / 1e6))
    if rank == 0: log("workload ready:", facts)
    return dict(ri=ri, tags=tags, layout=0, reads=reads, name=name, facts=facts)

def cpu_oracle_rate(o, reads_np, L, min_len, min_occ, target_s, threads=0, count_work=True):
    threads = threads or host_threads()
    """Time the oracle (all host threads) on a bounded sample sized for about target_s seconds. -> dict"""
    n_avail = len(reads_np) // L
    probe = min(n_avail, 20000)
    off = np.arange(probe + 1, dtype=np.uint64) * np.uint64(L)
    dt = max(1e-6, o.find_mems(reads_np[: probe * L], off, min_len, min_occ, threads=threads, text=False)["seconds"])
    n = int(min(n_avail, max(probe, probe * target_s / dt)))
    off = np.arange(n + 1, dtype=np.uint64) * np.uint64(L)
    R = o.find_mems(reads_np[: n * L], off, min_len, min_occ, threads=threads, count_work=count_work, text=False); dt = R["seconds"]
    return dict(reads=n, seconds=dt, rate=n / dt, result=R)

def run_reference(args, rank, world):
    """--impl reference: the CPU restatement of the reference's find_mems (oracle/), all host threads, rank 0 only."""
    if rank != 0:
        return
    import torch, orc
    w = make_workload(args, 0, 0)
    o = orc.Index(w["ri"], w["tags"], w["layout"])
    L = args.read_len
    reads_np = w["reads"][0].reshape(-1).cpu().numpy()
    cores = host_threads()
    per_step_s = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    probe = cpu_oracle_rate(o, reads_np, L, args.min_len, args.min_occ, per_step_s, count_work=False)
    n = probe["reads"]; off = np.arange(n + 1, dtype=np.uint64) * np.uint64(L)
    for _ in range(max(0, args.warmup - 1)):
        o.find_mems(reads_np[: n * L], off, args.min_len, args.min_occ, threads=cores, text=False)
    t = 0.0
    for _ in range(args.steps):
        t += o.find_mems(reads_np[: n * L], off, args.min_len, args.min_occ, threads=cores, text=False)["seconds"]
    rate = n * args.steps / t
    sample = "%d reads per step (bounded sample of the %d-read batch), %d OpenMP threads" % (n, args.reads_per_step, cores)
    line = {"impl": "reference", "metric": "find_mems reads/s (150bp, MEM + tag query)", "value": rate, "unit": "reads/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": {"workload": w["name"], "facts": w["facts"], "reads_per_step": n, "read_len": L, "min_len": args.min_len, "min_occ": args.min_occ,
                       "note": "the reference's own find_mems is single-threaded and cannot be built here (SDSL/GBWT absent): this is the CPU restatement in oracle/, parallelised over reads with OpenMP"},
            "cpu_baseline": {"value": rate, "unit": "reads/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": rate, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit_json(line)

def main():
    capture_stdout()
    args = parse_args()
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank, world)
    import torch, torch.distributed as dist
    from __graft_entry__ import load_package
    capi = load_package().capi
    if capi.lib().pgi_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the find_mems path has no CPU fallback")
    torch.cuda.set_device(local)
    # One process per GPU: run on the CPU cores next to this GPU (NVML's affinity mask), so that pinned staging buffers are
    # allocated on the GPU's own NUMA node and stream synchronisations wake a near core (r01r, 8 GPUs: ranks 0-3 moved
    # their results at 31 ms per step, ranks 4-7 at 22 ms). The CPU legs widen it again (host_threads).
    orig_affinity = None
    if world > 1:
        try:
            import pynvml
            pynvml.nvmlInit()
            orig_affinity = os.sched_getaffinity(0)
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", ""); idx = local
            if vis and all(x.strip().isdigit() for x in vis.split(",")): idx = int(vis.split(",")[local])
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(idx))
            if not (os.sched_getaffinity(0) & orig_affinity): os.sched_setaffinity(0, orig_affinity)
        except Exception as e:
            log("cpu affinity not set:", e)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    def barrier():
        if world > 1: dist.barrier()
        torch.cuda.synchronize()
    w = make_workload(args, local, rank)
    t0 = time.time()
    ix = capi.Index(w["ri"], w["tags"], w["layout"], device=local)
    index_create_s = time.time() - t0                       # parse + flatten on the host + upload: excluded from every timed region (SURVEY §8 d)
    if rank == 0: log("index on device: %.1f MB in %.1fs" % (ix.info["device_bytes"] / 1e6, index_create_s), ix.info)
    L = args.read_len; n = args.reads_per_step; pool = w["reads"].shape[0]
    d_off = torch.arange(n + 1, dtype=torch.int64, device="cuda") * L
    stream = torch.cuda.current_stream()
    batch = capi.Batch(ix, stream=stream.cuda_stream)
    # pinned host copies of the read batches for the end-to-end leg
    h_reads = [torch.empty(n * L + 16, dtype=torch.uint8).pin_memory() for _ in range(pool)]
    for i in range(pool): h_reads[i][: n * L].copy_(w["reads"][i].reshape(-1))
    h_off = (torch.arange(n + 1, dtype=torch.int64) * L).pin_memory()
    d_reads = [torch.cat([w["reads"][i].reshape(-1), torch.zeros(16, dtype=torch.uint8, device="cuda")]) for i in range(pool)]
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
    if not os.environ.get("PGI_BENCH_NOSAMPLER"): sampler.start()          # diagnostic switch (tools/placement_probe.py)
    barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    k1_ms = 0.0; stage = np.zeros(8); ext_total = 0; launches = 0; mems_total = 0; tags_total = 0; seed_total = 0
    e0.record(stream); host_gap = []; t_prev = time.perf_counter()
    for i in range(args.steps):
        step_device(args.warmup + i)
        t_now = time.perf_counter(); host_gap.append(1e3 * (t_now - t_prev)); t_prev = t_now      # host wall time of the step (diagnostic)
        ms = batch.stage_ms(); stage += np.array(ms); k1_ms += ms[0]; launches += int(ms[6]) + 0
        r = batch.device_result(); ext_total += r.extensions; mems_total += r.n_mems; tags_total += r.n_tags; seed_total += r.seed_lookups
    e1.record(stream)
    barrier()
    dev_ms = e0.elapsed_time(e1)
    # ---- end-to-end timing: K calls of pgi_find_mems (pinned host buffers in, all result arrays out to pinned host memory),
    # issued by E2E_CALLERS host threads the way the batched find_mems driver does, so the copies of one call overlap the
    # kernels of another. Wall clock around all K calls (every call blocks until its result is in host memory).
    import threading
    def e2e_pass(k_calls, first):
        nxt = [first]; lock = threading.Lock(); acc = [0]
        def caller():
            while True:
                with lock:
                    i = nxt[0]
                    if i >= first + k_calls: return
                    nxt[0] += 1
                nr_, nm_, nt_ = step_e2e(i)
                with lock: acc[0] += (nr_ + 1) * 8 + nm_ * 32 + (nm_ + 1) * 8 + nt_ * 8 + nm_ * 8
        ths = [threading.Thread(target=caller) for _ in range(min(E2E_CALLERS, k_calls))]
        t0 = time.perf_counter(); [t.start() for t in ths]; [t.join() for t in ths]
        return 1e3 * (time.perf_counter() - t0), acc[0]
    e2e_pass(max(args.warmup, E2E_CALLERS), 0)
    barrier()
    e2e_ms, d2h = e2e_pass(args.steps, args.warmup)
    barrier()
    sampler.stop_flag = True
    if sampler.is_alive(): sampler.join(timeout=2)
    # ---- max over ranks
    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([launches, ext_total, mems_total, tags_total], dtype=torch.float64, device="cuda")
    per_rank = [[dev_ms / args.steps, e2e_ms / args.steps, max(host_gap)]]
    if world > 1:
        mine = torch.tensor(per_rank[0], dtype=torch.float64, device="cuda"); allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine); per_rank = [[round(float(x), 3) for x in r] for r in allr]
        dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    dev_ms, e2e_ms = float(t[0]), float(t[1])
    if rank == 0: log("per rank [device ms/step, e2e ms/step, slowest host step ms]:", per_rank, "rank 0 host ms per step:", [round(x, 2) for x in host_gap])
    if rank != 0:
        if world > 1: dist.destroy_process_group()
        return
    total_reads = n * args.steps * world
    value = total_reads / (dev_ms * 1e-3); e2e = total_reads / (e2e_ms * 1e-3)
    # ---- CPU oracle: parity on a sample + work counters + cpu_baseline (rank 0, not timed against the GPU)
    if orig_affinity:
        try: os.sched_setaffinity(0, orig_affinity)          # the CPU legs use every host core again
        except Exception: pass
    import orc
    o = orc.Index(w["ri"], w["tags"], w["layout"])
    reads_np = h_reads[0][: n * L].numpy()
    ns = min(n, 20000); offs = np.arange(ns + 1, dtype=np.uint64) * np.uint64(L)
    R = o.find_mems(reads_np[: ns * L], offs, args.min_len, args.min_occ, count_work=True, text=False)
    G = ix.find_mems(reads_np[: ns * L], offs, args.min_len, args.min_occ, text=False)
    parity = bool(G["n_mems"] == R["n_mems"] and np.array_equal(G["mems"], R["mems"]) and np.array_equal(G["tags"], R["tags"])
                  and np.array_equal(G["n_runs"], R["n_runs"]) and G["extensions"] == R["counters"]["E"])
    # ---- next row of the scope table, reported beside the headline, not part of it: the query_tags path
    # (whole-read count_encoded + tag query of the final range) on the same reads, device-resident, with its own parity check
    next_rows = {}
    try:
        qs = []
        for i in range(2 + 3):
            batch.set_device_reads(d_reads[i % pool].data_ptr(), d_off.data_ptr(), n, n * L)
            batch.run_query_tags()
            if i >= 2: qs.append((batch.stage_ms()[5], batch.device_result().extensions))
        QG = ix.query_tags(reads_np[: ns * L], offs, text=False); QR = o.query_tags(reads_np[: ns * L], offs, text=False, threads=host_threads())
        q_par = bool(np.array_equal(QG["mems"], QR["mems"]) and np.array_equal(QG["tags"], QR["tags"]) and np.array_equal(QG["n_runs"], QR["n_runs"]))
        q_ms = float(np.median([x[0] for x in qs]))
        next_rows["query_tags"] = {"reads_per_s": n / (q_ms * 1e-3), "ms_per_step": q_ms, "lf_steps_per_read": qs[-1][1] / n,
                                   "matched_fraction": float((QR["mems"][:, 3] > 0).mean()), "parity_sample": "%d reads bit-exact vs oracle: %s" % (ns, q_par),
                                   "cpu_port_reads_per_s": ns / QR["seconds"], "note": "150 bp reads with 1 % errors searched whole (right to left until the range is empty)"}
    except Exception as e:
        next_rows["query_tags"] = {"error": str(e)}
    # ---- text formatting (src/find_mems.cpp:115-139 stdout grammar): host-bound, reported separately (SURVEY §8 d)
    fmt = {}
    try:
        import ctypes
        r = ix.find_mems_raw(h_reads[0].data_ptr(), h_off.data_ptr(), n, args.min_len, args.min_occ)
        ln = ctypes.c_uint64(); t1 = time.perf_counter()
        ptxt = capi.lib().pgi_format_text(ctypes.byref(r), 1, 1, ctypes.byref(ln)); t_fmt = time.perf_counter() - t1
        capi.lib().pgi_free(ptxt); ix.result_free(r)
        fmt = {"reads_per_s": n / t_fmt, "text_bytes_per_step": int(ln.value), "threads": 1, "note": "pgi_format_text, one host thread; not inside value / e2e"}
    except Exception as e:
        fmt = {"error": str(e)}
    c = R["counters"]; d_per_e = c["D"] / max(1, c["E"])
    cpu = cpu_oracle_rate(o, reads_np, L, args.min_len, args.min_occ, args.cpu_seconds, count_work=False)
    cores = host_threads()
    # ---- roofline of the dominant kernel (fmd_mem_kernel): algorithmic bytes = 128 B x D
    peaks = {}
    try: peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception: pass
    peak = float(peaks.get("hbm_gbs", 6650.0)); peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s"
    ext_rank0 = ext_total
    k1_bytes_per_launch = 128.0 * d_per_e * ext_rank0 / args.steps
    k1_s = (k1_ms / args.steps) * 1e-3
    achieved = k1_bytes_per_launch / k1_s / 1e9
    # DRAM traffic of the same kernel on the same default workload, from the committed ncu --set full capture (per launch)
    traffic = None; traffic_src = None
    if args.workload == "synth" and args.ref_len == 50_000_000 and args.n_hap == 20 and n == 2_000_000 and L == 150:
        try:
            import glob, re
            f = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_fmd_mem_kernel_raw.txt")))[-1]
            vals = {}
            for ln in open(f):
                m = re.match(r"(dram__bytes_(?:read|write)\.sum)\s+(\S+)\s+(\S+)", ln)
                if m: vals[m.group(1)] = float(m.group(3)) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[m.group(2)]
            traffic = vals["dram__bytes_read.sum"] + vals["dram__bytes_write.sum"]; traffic_src = os.path.relpath(f, ROOT)
        except Exception:
            pass
    rnd = {}
    try:
        for b in (32, 64):
            rnd["random%d_gbs" % b] = round(capi.random_access_gbs(local, 8 << 30, b, 1 << 27, 8), 1)
    except Exception as e:
        rnd["error"] = str(e)
    line = {"metric": "find_mems reads/s (150bp, MEM + tag query)", "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int64", "data": "synthetic",
            "config": {"workload": w["name"], "facts": w["facts"], "reads_per_step_per_gpu": n, "read_len": L, "substitution_rate": args.sub,
                       "min_len": args.min_len, "min_occ": args.min_occ, "index": "replicated per GPU", "device_index_bytes": ix.info["device_bytes"], "seed_k": ix.info["seed_k"],
                       "l2_policy": "read batch (%.0f MB) and index are cycled over %d distinct batches; inputs larger than L2" % (n * L / 1e6, pool),
                       "parity_sample": "%d reads bit-exact vs oracle: %s" % (ns, parity),
                       "mems_per_read": float(cnt[2]) / total_reads, "tags_per_read": float(cnt[3]) / total_reads,
                       "extensions_per_read": float(cnt[1]) / total_reads},
            "e2e": {"value": e2e, "unit": "reads/s", "h2d_bytes_per_step": n * L + (n + 1) * 8, "d2h_bytes_per_step": d2h // args.steps, "ms_per_step": e2e_ms / args.steps, "concurrent_callers": E2E_CALLERS},
            "gpu_launches": int(cnt[0]),
            "per_rank_ms_per_step": {"device": [r[0] for r in per_rank], "e2e": [r[1] for r in per_rank]},
            "stage_ms_per_step": {k: round(float(v) / args.steps, 3) for k, v in zip(["fmd_mem_kernel", "mem_order", "tag_plan", "tag_decode_sort", "tag_unique_compact", "run_total"], stage[:6])},
            "roofline": {"bound": "hbm", "kernel": "fmd_mem_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": k1_bytes_per_launch,
                         "seed_lookups_per_launch": seed_total // args.steps, "executed_extensions_per_launch": (ext_rank0 - seed_total * ix.info["seed_k"]) // args.steps,
                         "traffic_gbs": (traffic / k1_s / 1e9) if traffic else None, "frac_traffic": (traffic / k1_s / 1e9 / peak) if traffic else None,
                         "bytes_model": "128 B x D, D = distinct reference 10-run blocks touched per extension (oracle: D/E = %.3f) x E = %d extensions per launch (the reference algorithm's count = the oracle's E; with seed_k > 0 the first seed_k extensions of a walk from the full interval are ONE 16-byte table lookup, so the kernel moves fewer bytes than the algorithm names and frac may exceed 1)" % (d_per_e, ext_rank0 // args.steps),
                         "kernel_ms_per_launch": k1_ms / args.steps, **rnd},
            "cpu_baseline": {"value": cpu["rate"], "unit": "reads/s", "cores": cores, "kind": "port",
                             "sample": "%d reads of the same batch in %.1f s, %d OpenMP threads (oracle/ restatement; the reference itself is single-threaded)" % (cpu["reads"], cpu["seconds"], cores)},
            "index_create_s": round(index_create_s, 2), "text_format": fmt,
            "next_rows": next_rows,
            "clocks": sampler.summary()}
    emit_json(line)
    if world > 1: dist.destroy_process_group()

if __name__ == "__main__":
    main()
