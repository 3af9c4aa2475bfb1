#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.log
timeout 900 python -m pytest tests/test_gpu_admm.py -m gpu -q -x --timeout 300 > gpurun_out/pytest_admm.log 2>&1; echo "pytest admm exit $?" >> gpurun_out/summary.log
tail -30 gpurun_out/pytest_admm.log
for w in admm_65536x16384_k128 ao_admm_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w exit $?" >> gpurun_out/summary.log
  cat gpurun_out/bench_$w.json; tail -3 gpurun_out/bench_$w.err
done
cat gpurun_out/summary.log
