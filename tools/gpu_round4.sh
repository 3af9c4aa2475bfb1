#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.log
for t in gemm gram objective; do
  timeout 300 python tools/selftest.py $t > gpurun_out/st_$t.log 2>&1; echo "$t exit $?" >> gpurun_out/summary.log
  grep -c OK gpurun_out/st_$t.log; grep -E "BAD|Error|error" gpurun_out/st_$t.log | head -5
done
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/summary.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python tools/selftest.py bench 65536 65536 128 gemm > gpurun_out/bench_gemm_hl.log 2>&1; cat gpurun_out/bench_gemm_hl.log | tail -8
timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/bench_headline.json 2> gpurun_out/bench_headline.err; echo "bench exit $?" >> gpurun_out/summary.log
cat gpurun_out/bench_headline.json; tail -3 gpurun_out/bench_headline.err
cat gpurun_out/summary.log
