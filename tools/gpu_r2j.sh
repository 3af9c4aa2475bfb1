#!/bin/bash
# Round 2, call J (1 GPU): AO-ADMM back on the fp64 apply (pipelined); parity diagnostics (trajectory vs evaluation error).
mkdir -p gpurun_out; S=gpurun_out/r2j_summary.log; : > $S
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -k "admm or smoke or l2n" > gpurun_out/r2j_pytest.log 2>&1; echo "pytest exit $?" >> $S
grep -E "passed|failed" gpurun_out/r2j_pytest.log | tail -3 >> $S
timeout 900 python tools/diag_parity.py > gpurun_out/r2j_diag.log 2>&1; echo "diag exit $?" >> $S
cat gpurun_out/r2j_diag.log | cut -c1-900 >> $S
for w in admm_65536x16384_k128 ao_admm_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2j_bench_$w.json 2> gpurun_out/r2j_bench_$w.err; echo "bench $w exit $?" >> $S
  python - gpurun_out/r2j_bench_$w.json <<'PY' >> $S 2>&1
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s', round(d['ms_per_step'], 3), 'ms', d['phase_ms'], d.get('objective_last'))
PY
done
cat $S
