#!/usr/bin/env python3
"""Per-source-line summary of an ncu report (needs -lineinfo + --import-source on).
usage: tools/ncu_lines.py report.ncu-rep [top_n]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
cur = None; hdr = None; out = []
def num(x):
    try: return int(x)
    except Exception: return None
for r in rows:
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) > 3 and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] != "" and num(r[0]) is not None:
        s, i, t = num(r[6]), num(r[7]), num(r[8])
        if None in (s, i, t): continue
        out.append((cur, int(r[0]), r[1].strip()[:100], s, i, t))
ts = sum(o[3] for o in out); ti = sum(o[4] for o in out); tt = sum(o[5] for o in out)
print("total: samples %d  warp-inst %d  thread-inst %d  avg threads/inst %.2f" % (ts, ti, tt, tt / max(1, ti)))
out.sort(key=lambda o: -o[3 if (len(sys.argv) > 3 and sys.argv[3] == "samples") else 4])
for o in out[:top]:
    print("%-20s %4d  smp %5.1f%%  inst %5.1f%%  thr %5.1f  %s" % (o[0], o[1], 100 * o[3] / max(1, ts), 100 * o[4] / max(1, ti), o[5] / max(1, o[4]), o[2]))
