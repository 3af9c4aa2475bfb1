#!/bin/bash
# Round 2, call F (1 GPU): kl_data with K-blocked dual_v / D: parity of the KL branch, bench lines, launch list.
mkdir -p gpurun_out; S=gpurun_out/r2f_summary.log; : > $S
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -k "kl_branch" > gpurun_out/r2f_pytest.log 2>&1; echo "pytest exit $?" >> $S
tail -4 gpurun_out/r2f_pytest.log >> $S
for w in admm_kl_65536x16384_k128 ao_admm_kl_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2f_bench_$w.json 2> gpurun_out/r2f_bench_$w.err; echo "bench $w exit $?" >> $S
  tail -2 gpurun_out/r2f_bench_$w.err >> $S
  python - gpurun_out/r2f_bench_$w.json <<'PY' >> $S 2>&1
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s', round(d['ms_per_step'], 3), 'ms', d['phase_ms'], d.get('objective_last'))
PY
done
CMD="python bench.py --workload admm_kl_65536x16384_k128 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:nmfb" -s 70 -c 200 --csv --log-file gpurun_out/r2f_launches_admm_kl.csv $CMD > gpurun_out/r2f_ncu_list.log 2>&1
echo "ncu list exit $?" >> $S
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:kl_data_kernel" -s 3 -c 1 -o gpurun_out/full_kl_data_r02f -f $CMD > gpurun_out/r2f_ncu_full.log 2>&1
echo "ncu full exit $?" >> $S
ncu -i gpurun_out/full_kl_data_r02f.ncu-rep --page raw --csv > gpurun_out/full_kl_data_r02f.raw.csv 2>/dev/null
cat $S
