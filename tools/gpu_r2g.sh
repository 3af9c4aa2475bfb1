#!/bin/bash
# Round 2, call G (1 GPU): AO-ADMM inner iterations in split precision (fp64 base once, fp32 per iteration); kl_data Newton sqrt.
mkdir -p gpurun_out; S=gpurun_out/r2g_summary.log; : > $S
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -k "admm or kl_branch or smoke or l2n" > gpurun_out/r2g_pytest.log 2>&1; echo "pytest exit $?" >> $S
grep -E "relObj=|passed|failed" gpurun_out/r2g_pytest.log | tail -12 >> $S
for w in ao_admm_65536x16384_k128 admm_kl_65536x16384_k128 ao_admm_kl_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2g_bench_$w.json 2> gpurun_out/r2g_bench_$w.err; echo "bench $w exit $?" >> $S
  tail -2 gpurun_out/r2g_bench_$w.err >> $S
  python - gpurun_out/r2g_bench_$w.json <<'PY' >> $S 2>&1
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s', round(d['ms_per_step'], 3), 'ms', d['phase_ms'], d.get('objective_last'))
PY
done
CMD="python bench.py --workload ao_admm_65536x16384_k128 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:nmfb" -s 80 -c 200 --csv --log-file gpurun_out/r2g_launches_ao_admm.csv $CMD > gpurun_out/r2g_ncu_list.log 2>&1
echo "ncu list exit $?" >> $S
cat $S
