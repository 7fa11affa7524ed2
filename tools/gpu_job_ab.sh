#!/bin/bash
# A/B of library variants (tools/build_variant.sh): the headline bench with each of build_variants/libtb_<name>.so; usage: tools/gpu_job_ab.sh <name>...
mkdir -p gpurun_out
for v in base "$@"; do
  if [ $v = base ]; then unset TB_LIB_PATH; else export TB_LIB_PATH=$PWD/build_variants/libtb_$v.so; fi
  timeout 300 python bench.py --no-cpu-baseline --steps 4 > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err
  python -c "import json; d=json.load(open('gpurun_out/ab_$v.json')); k=d['kernels_ms_per_step']; print('$v', round(d['ms_per_step'],2), round(d['e2e']['ms_per_step'],2), {a: k[a] for a in list(k)[:12]})"
done
