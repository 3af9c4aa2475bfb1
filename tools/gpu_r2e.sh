#!/bin/bash
# Round 2, call E (1 GPU): k x k work on a side stream next to the numerator GEMM; kl_data with loads in flight.
mkdir -p gpurun_out; S=gpurun_out/r2e_summary.log; : > $S
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -k "admm or kl_branch or smoke or l2n" > gpurun_out/r2e_pytest.log 2>&1; echo "pytest exit $?" >> $S
tail -4 gpurun_out/r2e_pytest.log >> $S
for w in admm_65536x16384_k128 ao_admm_65536x16384_k128 admm_kl_65536x16384_k128 ao_admm_kl_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2e_bench_$w.json 2> gpurun_out/r2e_bench_$w.err; echo "bench $w exit $?" >> $S
  tail -2 gpurun_out/r2e_bench_$w.err >> $S
  python - gpurun_out/r2e_bench_$w.json <<'PY' >> $S 2>&1
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s', round(d['ms_per_step'], 3), 'ms', d['phase_ms'], d.get('objective_last'))
PY
done
cat $S
