mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 4 --no-cpu-baseline > gpurun_out/r02_bench_n4_auto.json 2> gpurun_out/r02_bench_n4_auto.err; echo rc=$?
python -c "import json; d=json.load(open('gpurun_out/r02_bench_n4_auto.json')); print(d['value'], d['ms_per_step'], d['e2e'])"
tail -3 gpurun_out/r02_bench_n4_auto.err
