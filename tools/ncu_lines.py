#!/usr/bin/env python3
"""Summarise an .ncu-rep: headline metrics + instructions / stall samples per source line.
usage: tools/ncu_lines.py report.ncu-rep [top_n]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, vals = rows[0], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit", "sm__warps_active.avg.per_cycle_active",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sass__inst_executed_local", "launch__grid_size"]
for i, h in enumerate(hdr):
    if any(h.startswith(w) for w in want) and "per_second" not in h and "pct_of_peak" not in h.replace("gpu__dram_throughput.avg.pct_of_peak", "").replace("smsp__issue_active.avg.pct_of_peak", ""):
        print("%-75s %s %s" % (h, vals[i], rows[1][i]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
cur = hdr = None
out = collections.defaultdict(lambda: [0, 0, 0])
for r in csv.reader(io.StringIO(src)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
        ii, si, ti = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
    elif hdr and r[0].isdigit():
        def f(x):
            try:
                return float(x)
            except ValueError:
                return 0.0
        k = (cur, int(r[0]), r[1].strip()[:80])
        out[k][0] += f(r[ii]); out[k][1] += f(r[si]); out[k][2] += f(r[ti])
tot = sum(v[0] for v in out.values()) or 1
tots = sum(v[1] for v in out.values()) or 1
byfile = collections.defaultdict(lambda: [0, 0])
for k, v in out.items():
    byfile[k[0]][0] += v[0]; byfile[k[0]][1] += v[1]
print()
for f_, v in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print("%-28s %5.1f%% inst %5.1f%% samples" % (f_, 100 * v[0] / tot, 100 * v[1] / tots))
print()
for k, v in sorted(out.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% inst %5.1f%% smp thr/inst %4.1f  %s:%d  %s" % (100 * v[0] / tot, 100 * v[1] / tots, v[2] / max(v[0], 1), k[0], k[1], k[2]))
print("\n-- by stall samples")
for k, v in sorted(out.items(), key=lambda kv: -kv[1][1])[:15]:
    print("%5.1f%% smp %5.1f%% inst  %s:%d  %s" % (100 * v[1] / tots, 100 * v[0] / tot, k[0], k[1], k[2]))
