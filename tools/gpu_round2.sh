#!/bin/bash
# Round-1 measurement pass: GPU tests, headline bench, ncu launch list + full capture of the panel GEMM.
mkdir -p gpurun_out
rm -f gpurun_out/summary.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/summary.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 30 --warmup 3 > gpurun_out/bench_headline.json 2> gpurun_out/bench_headline.err; echo "bench exit $?" >> gpurun_out/summary.log
cat gpurun_out/bench_headline.json
timeout 600 python bench.py --workload mur_eu_32768x32768_k64 --steps 50 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench_c2 exit $?" >> gpurun_out/summary.log
cat gpurun_out/bench_c2.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench_ref exit $?" >> gpurun_out/summary.log
cat gpurun_out/bench_ref.json
K='regex:panel_gemm|residual_kernel|mur_eu_update|gram_reduce|sum_partials'
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?" >> gpurun_out/summary.log
timeout 600 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:panel_gemm -s 6 -c 2 -o gpurun_out/prof_panel_gemm -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?" >> gpurun_out/summary.log
cat gpurun_out/summary.log
