#!/usr/bin/env python3
"""Summarise an .ncu-rep by source file: share of executed warp instructions and average active
threads per instruction, per kernel -- shows which part of a fused kernel body the time goes to.
usage: tools/ncu_regions.py report.ncu-rep [top_lines]"""
import collections
import csv
import subprocess
import sys
import os
sys.path.insert(0, "/root/repo/tools" if not os.path.exists(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_names.py")) else os.path.dirname(os.path.abspath(__file__)))
from ncu_names import kname

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 12
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
fn = fpath = hdr = None
agg = collections.defaultdict(lambda: collections.defaultdict(lambda: [0, 0]))
lines = collections.defaultdict(lambda: collections.defaultdict(lambda: [0, 0, ""]))
for r in csv.reader(src.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        fpath = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        fn = kname(r[1])
        continue
    if r[0] == "Line No":
        hdr = r
        ie = hdr.index("Instructions Executed")
        ti = hdr.index("Thread Instructions Executed")
        so = hdr.index("Source")
        continue
    if hdr and fn and r[0].isdigit() and r[2] == "-" and r[ie].isdigit():
        agg[fn][fpath][0] += int(r[ie])
        agg[fn][fpath][1] += int(r[ti])
        e = lines[fn][(fpath, int(r[0]))]
        e[0] += int(r[ie]); e[1] += int(r[ti]); e[2] = r[so].strip()[:110]
for fn, d in agg.items():
    tot = sum(v[0] for v in d.values())
    print("== %s  %.3g warp instructions" % (fn, tot))
    for k, v in sorted(d.items(), key=lambda kv: -kv[1][0]):
        print("  %5.1f%%  thr %4.1f  %s" % (100 * v[0] / tot, v[1] / max(v[0], 1), k))
    print("  -- hottest lines")
    for (f, ln), v in sorted(lines[fn].items(), key=lambda kv: -kv[1][0])[:top]:
        print("  %5.1f%%  thr %4.1f  %s:%d  %s" % (100 * v[0] / tot, v[1] / max(v[0], 1), f, ln, v[2]))
