#!/bin/bash
# First GPU contact: kernel self-tests (one process each), pytest -m gpu, micro-benchmarks.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
for t in gemm gram update objective mur; do
  timeout 300 python tools/selftest.py $t > gpurun_out/st_$t.log 2>&1
  echo "$t exit $?" >> gpurun_out/summary.log
  tail -3 gpurun_out/st_$t.log
done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/summary.log
timeout 600 python tools/selftest.py bench 32768 32768 64 all > gpurun_out/bench_c2.log 2>&1; echo "bench_c2 exit $?" >> gpurun_out/summary.log
timeout 900 python tools/selftest.py bench 65536 65536 128 all > gpurun_out/bench_hl.log 2>&1; echo "bench_hl exit $?" >> gpurun_out/summary.log
cat gpurun_out/summary.log
tail -20 gpurun_out/bench_c2.log gpurun_out/bench_hl.log
