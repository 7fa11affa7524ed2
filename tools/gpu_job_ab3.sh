#!/bin/bash
# A/B of library variants on BASELINE config 3 (long documents); usage: tools/gpu_job_ab3.sh <name>...
mkdir -p gpurun_out
for v in base "$@"; do
  if [ $v = base ]; then unset TB_LIB_PATH; else export TB_LIB_PATH=$PWD/build_variants/libtb_$v.so; fi
  timeout 400 python bench.py --config 3 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ab3_$v.json 2> gpurun_out/ab3_$v.err
  python -c "import json; d=json.load(open('gpurun_out/ab3_$v.json')); k=d['kernels_ms_per_step']; print('$v', round(d['value']), round(d['text_gbs'],2), round(d['gpu_ms_per_step'],2), {a: k[a] for a in list(k)[:5]})" || tail -3 gpurun_out/ab3_$v.err
done
unset TB_LIB_PATH
for opt in "--chunk-ramp 0" "--chunk-ramp 0 --chunk-mb 2048" "--chunk-ramp 0 --chunk-mb 512"; do
  timeout 400 python bench.py --config 3 --steps 2 --warmup 1 --no-cpu-baseline $opt > gpurun_out/ab3_opt.json 2> gpurun_out/ab3_opt.err
  python -c "import json; d=json.load(open('gpurun_out/ab3_opt.json')); k=d['kernels_ms_per_step']; print('$opt', round(d['value']), round(d['text_gbs'],2), round(d['gpu_ms_per_step'],2), {a: k[a] for a in list(k)[:5]})" || tail -3 gpurun_out/ab3_opt.err
done
