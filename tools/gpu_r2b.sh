#!/bin/bash
# Round 2, call B (1 GPU): the whole GPU suite after the kernel clean-up (incl. the mid / full-size goldens) + benches.
mkdir -p gpurun_out; S=gpurun_out/r2b_summary.log; : > $S
timeout 2400 python -m pytest tests -m gpu -q -rA --timeout 900 --durations=15 > gpurun_out/r2b_pytest.log 2>&1; echo "pytest exit $?" >> $S
grep -E "passed|failed|error" gpurun_out/r2b_pytest.log | tail -3 >> $S
grep -E "^(FAILED|ERROR)" gpurun_out/r2b_pytest.log | head -20 >> $S
grep -E "relObj=" gpurun_out/r2b_pytest.log >> $S
for w in mur_eu_65536x65536_k128 mur_kl_20000x50000_k128 admm_65536x16384_k128 ao_admm_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu > gpurun_out/r2b_bench_$w.json 2> gpurun_out/r2b_bench_$w.err; echo "bench $w exit $?" >> $S
  python - $w <<'PY' >> gpurun_out/r2b_summary.log
import json, sys
for ln in open('gpurun_out/r2b_bench_%s.json' % sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s e2e', d.get('e2e', {}) and d['e2e'].get('value'), 'phase_ms', d.get('phase_ms'))
PY
done
cat $S
