#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.log
timeout 600 python -m pytest tests/test_gpu_kl.py -m gpu -q --timeout 120 > gpurun_out/pytest_kl.log 2>&1; echo "pytest kl exit $?" >> gpurun_out/summary.log
grep -E "^E  .*assert|passed|failed|Error|timeout" gpurun_out/pytest_kl.log | head -30
timeout 600 python bench.py --workload mur_kl_20000x50000_k128 --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_kl.json 2> gpurun_out/bench_kl.err; echo "bench kl exit $?" >> gpurun_out/summary.log
cat gpurun_out/bench_kl.json | cut -c1-1500; tail -5 gpurun_out/bench_kl.err
cat gpurun_out/summary.log
