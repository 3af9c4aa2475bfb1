#!/bin/bash
# Round 2, call K (1 GPU): Gram of the k x k systems in fp64 -- component accuracy, parity diagnostics, ADMM family tests + benches.
mkdir -p gpurun_out; S=gpurun_out/r2k_summary.log; : > $S
timeout 600 python tools/diag_admm_step.py 8 > gpurun_out/r2k_diag_step.log 2>&1; echo "diag step exit $?" >> $S; cat gpurun_out/r2k_diag_step.log >> $S
timeout 900 python tools/diag_parity.py > gpurun_out/r2k_diag.log 2>&1; echo "diag exit $?" >> $S
grep -E "max relObj|after" gpurun_out/r2k_diag.log | cut -c1-330 >> $S
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -k "admm or smoke or l2n or kl_branch or mid_ or full_c4" > gpurun_out/r2k_pytest.log 2>&1; echo "pytest exit $?" >> $S
grep -E "relObj=|passed|failed" gpurun_out/r2k_pytest.log | tail -12 >> $S
for w in admm_65536x16384_k128 ao_admm_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2k_bench_$w.json 2> gpurun_out/r2k_bench_$w.err; echo "bench $w exit $?" >> $S
  python - gpurun_out/r2k_bench_$w.json <<'PY' >> $S 2>&1
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s', round(d['ms_per_step'], 3), 'ms', d['phase_ms'], d.get('objective_last'))
PY
done
cat $S
