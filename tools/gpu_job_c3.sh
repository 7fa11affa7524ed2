mkdir -p gpurun_out
for opt in "" "--chunk-mb 512" "--chunk-mb 768"; do
  timeout 400 python bench.py --config 3 --steps 2 --warmup 1 --no-cpu-baseline $opt > gpurun_out/c3_opt.json 2> gpurun_out/c3_opt.err
  python -c "import json; d=json.load(open('gpurun_out/c3_opt.json')); k=d['kernels_ms_per_step']; print('[$opt]', round(d['value']), round(d['text_gbs'],2), round(d['gpu_ms_per_step'],2), {a: k[a] for a in list(k)[:5]})" || tail -3 gpurun_out/c3_opt.err
done
timeout 300 python bench.py --no-cpu-baseline --e2e-api call > gpurun_out/c3_head.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/c3_head.json')); print('headline', round(d['ms_per_step'],2), round(d['e2e']['ms_per_step'],2))"
