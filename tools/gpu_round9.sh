#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.log
timeout 300 python tools/selftest.py objective > gpurun_out/st_objective.log 2>&1; echo "objective exit $?" >> gpurun_out/summary.log
grep -c "OK " gpurun_out/st_objective.log
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/summary.log
tail -8 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 30 --warmup 3 > gpurun_out/bench_headline.json 2> gpurun_out/bench_headline.err; echo "bench exit $?" >> gpurun_out/summary.log
cat gpurun_out/bench_headline.json; tail -3 gpurun_out/bench_headline.err
for w in mur_eu_32768x32768_k64 admm_65536x16384_k128 ao_admm_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w exit $?" >> gpurun_out/summary.log
  python -c "
import json,sys
d=json.load(open('gpurun_out/bench_$w.json')); print('$w', round(d['value'],1),'it/s', d['phase_ms'])"
done
cat gpurun_out/summary.log
