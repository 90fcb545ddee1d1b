#!/usr/bin/env python3
"""Summarise an `ncu --set full` capture of fmd_mem_kernel: writes profiles/<tag>_fmd_mem_kernel_raw.txt (selected raw
metrics, per launch) and records the DRAM bytes per launch in profiles/k1_traffic.json keyed by the SHA-256 of the files that
define fmd_mem_kernel and its data layout (csrc/pgi_kernels.cuh, pgi_core.cuh, pgi_layout.h; bench.py
quotes `roofline.traffic` only from an entry whose hash matches the kernel it is running).
usage: tools/ncu_k1_summary.py report.ncu-rep tag workload reads_per_step [nominal=1]"""
import csv, hashlib, io, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, tag, workload, reads = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]); nominal = (sys.argv[5] != "0") if len(sys.argv) > 5 else True
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt))); hdr, units, data = rows[0], rows[1], rows[2:]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__grid_size", "launch__block_size", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum", "lts__t_requests_srcunit_tex_op_read.sum", "dram__sectors_read.sum",
        "sm__cycles_active.avg", "gpc__cycles_elapsed.max", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed"]
lines = ["# ncu --set full --clock-control none, kernel fmd_mem_kernel, %d launch(es); workload %s, %d reads per launch" % (len(data), workload, reads)]
vals = {}
for k in want:
    if k in hdr:
        i = hdr.index(k); vals[k] = [r[i] for r in data]; lines.append("%-90s %-12s %s" % (k, units[i], "  ".join(vals[k])))
open(os.path.join(ROOT, "profiles", tag + "_fmd_mem_kernel_raw.txt"), "w").write("\n".join(lines) + "\n")
def to_bytes(k):
    i = hdr.index(k); mul = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[units[i]]
    return sum(float(r[i]) for r in data) / len(data) * mul
hh = hashlib.sha256(); d = os.path.join(ROOT, "pangenome-index_b200", "csrc")
for f in sorted(os.listdir(d)):
    if f in ("pgi_kernels.cuh", "pgi_core.cuh", "pgi_layout.h"): hh.update(open(os.path.join(d, f), "rb").read())
p = os.path.join(ROOT, "profiles", "k1_traffic.json")
tj = json.load(open(p)) if os.path.exists(p) else []
tj = [e for e in tj if not (e["csrc_sha256"] == hh.hexdigest() and e["workload"] == workload and e["reads_per_step"] == reads)]
tj.append({"csrc_sha256": hh.hexdigest(), "workload": workload, "nominal": nominal, "reads_per_step": reads, "dram_bytes_per_launch": to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum"),
           "kernel_ms_under_ncu": sum(float(x) for x in vals["gpu__time_duration.sum"]) / len(data), "source": "profiles/%s_fmd_mem_kernel_raw.txt" % tag})
json.dump(tj, open(p, "w"), indent=1)
print("\n".join(lines))
