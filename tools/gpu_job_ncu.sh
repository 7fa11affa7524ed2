#!/bin/bash
# one full ncu capture of selected kernels (first chunk of a 200k-document run); usage: tools/gpu_job_ncu.sh <regex> <count> <outname>
mkdir -p gpurun_out
timeout 300 python bench.py --docs 200000 --steps 1 --warmup 1 --no-cpu-baseline --e2e-api call > gpurun_out/$3_plain.json 2> gpurun_out/$3_plain.err && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$1" -c $2 \
      -o gpurun_out/$3 python bench.py --docs 200000 --steps 1 --warmup 1 --no-cpu-baseline --e2e-api call > gpurun_out/$3_ncu.log 2>&1
echo "capture rc=$?"
