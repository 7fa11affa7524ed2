mkdir -p gpurun_out
for cfg in "4 32768" "4 131072" "3 65536" "5 65536" "6 43478" "8 32768"; do set -- $cfg; TB_BENCH_DEBUG=1 timeout 300 python bench.py --no-cpu-baseline --e2e-api runner --pipelines $1 --stream-group-docs $2 --steps 4 > gpurun_out/sweep.json 2> gpurun_out/sweep.err; echo "pipelines $1 group $2: $(grep '\[bench\]' gpurun_out/sweep.err)"; done
