#!/usr/bin/env python3
"""Which source lines own the hot code: executed SASS instructions per source line (static count) for one
kernel of an .ncu-rep.  usage: tools/ncu_static.py report.ncu-rep kernel_substring [top]"""
import collections
import csv
import subprocess
import sys

rep, want = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
fn = fpath = None
ie = None
cur = None
agg = collections.defaultdict(lambda: [0, 0])
for r in csv.reader(src.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        fpath = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        fn = r[1]
        continue
    if r[0] == "Line No":
        ie = r.index("Instructions Executed")
        continue
    if ie is None or want not in (fn or ""):
        continue
    if r[0].isdigit():
        cur = (fpath, int(r[0]), r[1].strip()[:90])
    elif r[0] == "" and cur and len(r) > ie and r[ie].isdigit() and int(r[ie]) > 0:
        agg[cur][0] += 1
        agg[cur][1] += int(r[ie])
tot = sum(v[1] for v in agg.values()) or 1
byfile = collections.defaultdict(int)
for k, v in agg.items():
    byfile[k[0]] += v[0]
print("executed SASS instructions (static):", sum(v[0] for v in agg.values()), dict(byfile))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%4d sass  %5.1f%% dyn  %s:%d  %s" % (v[0], 100 * v[1] / tot, k[0], k[1], k[2]))
