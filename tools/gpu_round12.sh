#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.log
timeout 900 python -m pytest tests/test_gpu_admm.py -m gpu -q --timeout 300 > gpurun_out/pytest_admm.log 2>&1; echo "pytest admm exit $?" >> gpurun_out/summary.log
grep -E "^E  .*assert [a-z0-9.(]|passed|failed|Error|timeout|FAILED" gpurun_out/pytest_admm.log | cut -c1-200 | head -30
python tools/diag_ao.py 2>&1 | awk 'NR<=14 || /iter1|admm_iter/' | cut -c1-120
for w in admm_65536x16384_k128 ao_admm_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w exit $?" >> gpurun_out/summary.log
  python -c "
import json
d=json.load(open('gpurun_out/bench_$w.json')); print('$w', round(d['value'],1),'it/s', d['phase_ms'])" || tail -5 gpurun_out/bench_$w.err
done
cat gpurun_out/summary.log
