#!/bin/bash
# Round 2, call M (1 GPU): KP = 256 kernels at 1/16 of config 5; headline + KL benches with launch lists for profiles/.
mkdir -p gpurun_out; S=gpurun_out/r2m_summary.log; : > $S
line() { python - "$1" <<'PY' >> gpurun_out/r2m_summary.log 2>&1
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        r = d.get('roofline') or {}
        print('  value', round(d['value'], 1), 'it/s', round(d['ms_per_step'], 3), 'ms', d['phase_ms'], 'roof', r.get('bound'), round(r.get('frac') or 0, 3), 'obj', d.get('objective_last'))
PY
}
for w in mur_eu_65536x32768_k256 mur_eu_32768x32768_k64 mur_kl_20000x50000_k128; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2m_bench_$w.json 2> gpurun_out/r2m_bench_$w.err; echo "bench $w exit $?" >> $S
  tail -2 gpurun_out/r2m_bench_$w.err >> $S; line gpurun_out/r2m_bench_$w.json
done
timeout 900 python -m pytest tests -m gpu -q -x --timeout 900 -k "shapes_vs_oracle or tmem_operand or wide_tile" > gpurun_out/r2m_pytest.log 2>&1; echo "pytest exit $?" >> $S
tail -3 gpurun_out/r2m_pytest.log >> $S
cat $S
