#!/bin/bash
# usage: tools/gpu_job_multi.sh N  -- headline bench on N GPUs in both content modes, then config 4
N=$1
mkdir -p gpurun_out
RUN="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N"
$RUN --no-cpu-baseline > gpurun_out/r02_bench_n${N}_runs.json 2> gpurun_out/r02_bench_n${N}_runs.err; echo "runs rc=$?"
if [ "$N" -lt 8 ]; then $RUN --no-cpu-baseline --content bytes > gpurun_out/r02_bench_n${N}_bytes.json 2> gpurun_out/r02_bench_n${N}_bytes.err; echo "bytes rc=$?"; fi
$RUN --config 4 --no-cpu-baseline > gpurun_out/r02_config4_n${N}.json 2> gpurun_out/r02_config4_n${N}.err; echo "config4 rc=$?"
for f in runs bytes; do [ -s gpurun_out/r02_bench_n${N}_$f.json ] && python -c "import json; d=json.load(open('gpurun_out/r02_bench_n${N}_$f.json')); print('$f', d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'])"; done
python -c "import json; d=json.load(open('gpurun_out/r02_config4_n${N}.json')); print('config4', d['value'], d['ms_per_step'], d.get('gpu_ms_per_step'))"
