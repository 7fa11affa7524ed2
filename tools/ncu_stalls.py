#!/usr/bin/env python3
"""Warp stall reasons per kernel from an .ncu-rep (stalled warps per issue-active cycle).
usage: tools/ncu_stalls.py report.ncu-rep"""
import csv
import subprocess
import sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")].split("::")[-1][:14]
    out = []
    for i, h in enumerate(hdr):
        if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
            try:
                v = float(r[i])
            except ValueError:
                continue
            out.append((v, h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
    out.sort(reverse=True)
    def g(m):
        return r[hdr.index(m)] if m in hdr else "?"
    print(name, "issue_active%", g("smsp__issue_active.avg.pct_of_peak_sustained_active"), "warps_active%", g("sm__warps_active.avg.pct_of_peak_sustained_active"))
    print("   ", "  ".join("%s %.2f" % (n, v) for v, n in out[:7]))
