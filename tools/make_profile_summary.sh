#!/bin/bash
# Rebuild profiles/r02_* from the captures in gpurun_out/ (launch list, full ncu report, bench lines).
# usage: tools/make_profile_summary.sh <full-report.ncu-rep>
set -e
REP=${1:-gpurun_out/r02_full_top4.ncu-rep}
ncu -i "$REP" --page raw --csv 2>/dev/null | python -c "
import csv,sys,json
rows=list(csv.reader(sys.stdin)); hdr=rows[0]; units=rows[1]; out={}
for r in rows[2:]:
    name=r[hdr.index('Kernel Name')].split('(')[0].split('::')[-1]
    # the names the library's event timers (and bench.py) use for template instantiations
    for k,v in {'k_gr_rep_w<0':'k_gr_rep_w256','k_gr_rep_w<1':'k_gr_rep_w512','k_gr_rep_c<2>':'k_gr_rep_c1k','k_gr_rep_c<3>':'k_gr_rep_c2k','k_gr_rep_c<4>':'k_gr_rep_c4k','k_gr_rep_c<5>':'k_gr_rep_c8k'}.items():
        if name.startswith(k): name=v
    name=name.split('<')[0]
    def val(m):
        i=hdr.index(m); v=float(r[i].replace(',','')); return v*{'byte':1,'Kbyte':1e3,'Mbyte':1e6,'Gbyte':1e9}[units[i]]
    out[name]=val('dram__bytes_read.sum')+val('dram__bytes_write.sum')
# the captured launch is the first chunk of the device-resident run: with the default plan that is 256 MiB of text
# (268.4 MB) and ~75.9k documents of 24 bytes of offsets / results each
out['_captured_launch_algorithmic_bytes']=float(sys.argv[1]) if len(sys.argv)>1 else 270.26e6
json.dump(out, open('profiles/roofline_traffic.json','w'), indent=1)
"
{
echo "# Round 2 -- ncu evidence for the bench command (python bench.py, 1M docs per step, chunks ramping up to 1 GiB)"
echo "# capture: ncu --set full --clock-control none --import-source on, first launch of each listed kernel (the first chunk of the device-resident run: 256 MiB, ~76k docs)"
echo "# launch list: profiles/r02_launches.csv (ncu --metrics gpu__time_duration.sum --clock-control none, serialised, cold caches)"
echo
echo "## kernel share of the step: ncu launch list vs CUDA events inside bench.py (profiles/r02_bench_n1_1Mdocs.json)"
python - <<'PY'
import csv, collections, json
rows=list(csv.reader(open('gpurun_out/r02_launches.csv')))
hi=[i for i,r in enumerate(rows) if r and r[0]=="ID"][0]
hdr=rows[hi]; kn=hdr.index("Kernel Name"); mv=hdr.index("Metric Value"); mu=hdr.index("Metric Unit")
# ncu prints template instantiations (k_gq<0>, k_gr_rep_w<0, 32>, k_gr_rep_c<3>); the library's event timers use the
# names below
TIER={"k_gr_rep_w<0":"k_gr_rep_w256","k_gr_rep_w<1":"k_gr_rep_w512","k_gr_rep_c<2>":"k_gr_rep_c1k","k_gr_rep_c<3>":"k_gr_rep_c2k","k_gr_rep_c<4>":"k_gr_rep_c4k","k_gr_rep_c<5>":"k_gr_rep_c8k"}
def canon(n):
    for k,v in TIER.items():
        if n.startswith(k): return v
    return n.split("<")[0]
agg=collections.defaultdict(lambda:[0,0.0])
for r in rows[hi+1:]:
    if len(r)<=mv: continue
    name=canon(r[kn].split("(")[0].split("::")[-1]); v=float(r[mv].replace(",","")); u=r[mu]
    agg[name][0]+=1; agg[name][1]+= v*1e3 if u=="us" else v*1e6 if u=="ms" else v
tot=sum(v[1] for v in agg.values())
live=json.load(open('gpurun_out/r02_bench_n1.json'))["kernels_ms_per_step"]; lt=sum(live.values())
print("%-18s %9s %12s %8s %10s"%("kernel","launches","ncu ms","ncu %","events %"))
for k,v in sorted(agg.items(), key=lambda kv:-kv[1][1]):
    print("%-18s %9d %12.3f %7.1f%% %9s"%(k,v[0],v[1]/1e6,100*v[1]/tot, ("%.1f%%"%(100*live[k]/lt)) if k in live else "-"))
print("(k_gr_rep_x runs beside the other tiers on a second stream in the real run; ncu serialises it)")
PY
echo
echo "## raw metrics, stall reasons, hot-code footprint"
python tools/ncu_hot.py "$REP" 2>/dev/null | grep "^k_"
python tools/ncu_stalls.py "$REP"
python tools/ncu_footprint.py "$REP"
echo
echo "## where the instructions go (per source file, hottest lines)"
python tools/ncu_regions.py "$REP" 14
} > profiles/r02_ncu_summary.txt 2>&1
cp gpurun_out/r02_launches.csv profiles/r02_launches.csv
cp gpurun_out/r02_bench_n1.json profiles/r02_bench_n1_1Mdocs.json
cp gpurun_out/r02_bench_reference.json profiles/r02_bench_reference_arm.json
cp gpurun_out/r02_gpu_tests.log profiles/r02_gpu_tests.log
[ -f gpurun_out/r02_bench_n1_with_fineweb.json ] && cp gpurun_out/r02_bench_n1_with_fineweb.json profiles/r02_bench_n1_with_fineweb.json
echo done
