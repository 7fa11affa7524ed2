#!/usr/bin/env python3
"""Hot instruction footprint per kernel from an .ncu-rep: how many SASS instructions (x 16 bytes) carry
90% / 99% of the executed warp instructions -- compare with the 32 KB L1.5 instruction cache.
usage: tools/ncu_footprint.py report.ncu-rep"""
import collections
import csv
import subprocess
import sys
import os
sys.path.insert(0, "/root/repo/tools" if not os.path.exists(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_names.py")) else os.path.dirname(os.path.abspath(__file__)))
from ncu_names import kname

src = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
fn = None
ie = None
stats = collections.defaultdict(list)
for r in csv.reader(src.splitlines()):
    if not r:
        continue
    if r[0] == "Kernel Name":
        fn = kname(r[1])
        continue
    if r[0] == "Address":
        ie = r.index("Instructions Executed")
        continue
    if fn and ie is not None and len(r) > ie and r[ie].isdigit():
        stats[fn].append(int(r[ie]))
for fn, v in stats.items():
    tot = sum(v)
    acc = 0
    k90 = k99 = None
    for i, x in enumerate(sorted(v, reverse=True)):
        acc += x
        if k90 is None and acc >= 0.9 * tot:
            k90 = i + 1
        if acc >= 0.99 * tot:
            k99 = i + 1
            break
    print("%-12s %6d SASS instructions (%4d KB), %6d ever executed; 90%% of the dynamic count in %5d (%3d KB), 99%% in %5d (%3d KB)"
          % (fn, len(v), len(v) * 16 // 1024, sum(1 for x in v if x), k90, k90 * 16 // 1024, k99, k99 * 16 // 1024))
