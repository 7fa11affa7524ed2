#!/bin/bash
# Round-2 evidence on one B200 (run under gpurun): the bench line of both arms, the ncu launch list of the bench command
# and one full capture of the six heaviest kernels (first chunk of a 200k-document run).  tools/make_profile_summary.sh
# turns the captures into profiles/r02_*.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gpu_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/r02_gpu_tests.log
timeout 300 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference > gpurun_out/r02_bench_reference.json 2>/dev/null; echo "reference rc=$?"
timeout 300 python bench.py --content bytes --no-cpu-baseline > gpurun_out/r02_bench_n1_bytes.json 2>/dev/null
timeout 300 python bench.py --order shipped --no-cpu-baseline > gpurun_out/r02_bench_n1_shipped_order.json 2>/dev/null
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --e2e-api call > gpurun_out/r02_plain.json 2> gpurun_out/r02_plain.err && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_launches.csv \
      python bench.py --steps 2 --warmup 1 --no-cpu-baseline --e2e-api call > gpurun_out/r02_ncu_run.log 2>&1
echo "launch list rc=$?"
timeout 300 python bench.py --docs 200000 --steps 1 --warmup 1 --no-cpu-baseline --e2e-api call > gpurun_out/r02_plain2.json 2> gpurun_out/r02_plain2.err && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_gr_rep_w|k_gr_rep_c|k_gr_scan|k_gq|k_c4_b|k_c4_a" -c 10 \
      -o gpurun_out/r02_full_top4 python bench.py --docs 200000 --steps 1 --warmup 1 --no-cpu-baseline --e2e-api call > gpurun_out/r02_ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out | tail -20
