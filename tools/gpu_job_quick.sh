mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "bulk or parity_shipped or runs or noalloc or repetition or tier" 2>&1 | tail -2
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r02q_bench_n1.json 2>gpurun_out/r02q_bench.err
python -c "import json; d=json.load(open('gpurun_out/r02q_bench_n1.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['kernels_ms_per_step'])"
timeout 600 python bench.py --config 3 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02q_config3.json 2>gpurun_out/r02q_c3.err
python -c "import json; d=json.load(open('gpurun_out/r02q_config3.json')); print('config3', d['value'], d['text_gbs'], d['kernels_ms_per_step'])"
