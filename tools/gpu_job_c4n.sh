mkdir -p gpurun_out
N=$1
RUN="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --config 4 --no-cpu-baseline"
$RUN > gpurun_out/r02_config4_n${N}_shipped.json 2> gpurun_out/r02_config4_n${N}.err; echo "shipped rc=$?"
$RUN --order baseline > gpurun_out/r02_config4_n${N}_baseline_order.json 2>> gpurun_out/r02_config4_n${N}.err; echo "baseline rc=$?"
for f in shipped baseline_order; do python -c "import json; d=json.load(open('gpurun_out/r02_config4_n${N}_$f.json')); print('$f', d['value'], d['text_gbs'], d['config']['runner'][:120])"; done
