mkdir -p gpurun_out
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02i_gpu_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r02i_gpu_tests.log
timeout 300 python bench.py > gpurun_out/r02i_bench_n1.json 2>gpurun_out/r02i_bench.err
python -c "import json; d=json.load(open('gpurun_out/r02i_bench_n1.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['parity']['mismatches'], d['kernels_ms_per_step'])"
for p in 2 3 4; do timeout 300 python bench.py --config 4 --no-cpu-baseline --pipelines $p > gpurun_out/r02i_config4_n1_p$p.json 2> gpurun_out/r02i_c4.err; python -c "import json; d=json.load(open('gpurun_out/r02i_config4_n1_p$p.json')); print($p, d['value'], d['ms_per_step'], d['gpu_ms_per_step'])"; done
timeout 600 python bench.py --config 3 --steps 2 --warmup 1 > gpurun_out/r02i_config3.json 2>gpurun_out/r02i_c3.err
python -c "import json; d=json.load(open('gpurun_out/r02i_config3.json')); print('config3', d['value'], d['text_gbs'], d.get('parity'), d['kernels_ms_per_step'])"
timeout 600 python bench.py --config 3 --steps 2 --warmup 1 --no-cpu-baseline --bulk-stage-tiles 8 > gpurun_out/r02i_config3_bulk8.json 2>/dev/null
python -c "import json; d=json.load(open('gpurun_out/r02i_config3_bulk8.json')); print('config3 bulk8', d['value'], d['text_gbs'], d['kernels_ms_per_step'])"
