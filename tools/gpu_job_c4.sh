mkdir -p gpurun_out
for p in 3 4 6; do timeout 300 python bench.py --config 4 --no-cpu-baseline --pipelines $p > gpurun_out/r02_config4_n1_p$p.json 2> gpurun_out/r02_c4.err; python -c "import json; d=json.load(open('gpurun_out/r02_config4_n1_p$p.json')); print($p, d['value'], d['ms_per_step'], d.get('gpu_ms_per_step'), d.get('text_gbs'))"; done
timeout 300 python bench.py --config 4 --no-cpu-baseline --pipelines 4 --order baseline > gpurun_out/r02_config4_n1_p4_baseline.json 2>> gpurun_out/r02_c4.err; python -c "import json; d=json.load(open('gpurun_out/r02_config4_n1_p4_baseline.json')); print('baseline order', d['value'], d.get('text_gbs'))"
tail -3 gpurun_out/r02_c4.err
