#!/bin/bash
# 1 GPU: tests + benches after the Cholesky / residual-pipeline changes.
mkdir -p gpurun_out; rm -f gpurun_out/summary.log
S=gpurun_out/summary.log
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $S
tail -5 gpurun_out/pytest_gpu.log
for w in mur_eu_65536x65536_k128 mur_eu_32768x32768_k64 mur_kl_20000x50000_k128 admm_65536x16384_k128 ao_admm_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w exit $?" >> $S
  python -c "
import json
d=json.load(open('gpurun_out/bench_$w.json')); print('$w', round(d['value'],1),'it/s', d['phase_ms'])" || tail -5 gpurun_out/bench_$w.err
done
cat $S
