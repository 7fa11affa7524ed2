#!/usr/bin/env python3
"""Summarise an .ncu-rep: per kernel headline metrics + hottest source lines (needs -lineinfo + --import-source on)."""
import collections, csv, subprocess, sys
import os
sys.path.insert(0, "/root/repo/tools" if not os.path.exists(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_names.py")) else os.path.dirname(os.path.abspath(__file__)))
from ncu_names import kname

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 18
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h = rows[0]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]
for r in rows[2:]:
    name = kname(r[h.index("Kernel Name")])
    print(name, {w.split(".")[0][-22:]: r[h.index(w)] for w in want if w in h})
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
fn = fpath = hdr = None
agg = collections.defaultdict(dict)
for r in csv.reader(src.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        fpath = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        fn = kname(r[1]); continue
    if r[0] == "Line No":
        hdr = r; ie = hdr.index("Instructions Executed"); si = hdr.index("# Samples"); ti = hdr.index("Thread Instructions Executed"); continue
    if hdr and fn and r[0].isdigit() and r[2] == "-" and r[ie].isdigit():
        agg[fn][(fpath, int(r[0]))] = [int(r[ie]), int(r[si] or 0), int(r[ti] or 0), r[1]]
for fn, d in agg.items():
    tot = sum(v[0] for v in d.values()); tots = sum(v[1] for v in d.values())
    print("=====", fn, "warp-instr %.3g" % tot, "samples", tots)
    for (f, l), v in sorted(d.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%5.1f%% smp%5.1f%% thr %4.1f  %s:%d  %s" % (100 * v[0] / tot, 100 * v[1] / max(tots, 1), v[2] / max(v[0], 1), f[:18], l, v[3].strip()[:100]))
