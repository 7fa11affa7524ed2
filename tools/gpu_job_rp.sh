mkdir -p gpurun_out
TB_LIB_PATH=$PWD/build_variants/libtb_prof.so python tools/rep_profile.py 200000 2>&1 | tail -9
timeout 300 python bench.py --no-cpu-baseline --e2e-api call > gpurun_out/rp_bench.json 2>/dev/null
python -c "import json; d=json.load(open('gpurun_out/rp_bench.json')); print(round(d['ms_per_step'],2), round(d['e2e']['ms_per_step'],2), d['kernels_ms_per_step'])"
