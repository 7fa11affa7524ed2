mkdir -p gpurun_out
timeout 300 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo rc=$?
python -c "import json; d=json.load(open('gpurun_out/r02_bench_n1.json')); print(round(d['ms_per_step'],2), d['e2e'], d['h2d_roofline']['frac'], d['parity'], d['roofline']['traffic'])"
timeout 300 python bench.py --order shipped --no-cpu-baseline > gpurun_out/r02_bench_n1_shipped_order.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/r02_bench_n1_shipped_order.json')); print('shipped', round(d['ms_per_step'],2), d['e2e']['value'], d['e2e']['api'][:40], d['e2e']['single_call'], d['e2e']['streamed'])"
timeout 300 python bench.py --content bytes --no-cpu-baseline > gpurun_out/r02_bench_n1_bytes.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/r02_bench_n1_bytes.json')); print('bytes', round(d['ms_per_step'],2), d['e2e']['value'], d['e2e']['api'][:40], d['e2e']['single_call'], d['e2e']['streamed'])"
