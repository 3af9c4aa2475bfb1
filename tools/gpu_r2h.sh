#!/bin/bash
# Round 2, call H (1 GPU): per-kernel times of the AO-ADMM and ADMM-KL steps, ncu --set full of admm_apply32 and kl_data.
mkdir -p gpurun_out; S=gpurun_out/r2h_summary.log; : > $S
timeout 600 python -m pytest tests -m gpu -q -x --timeout 900 -k "kl_branch" > gpurun_out/r2h_pytest.log 2>&1; echo "pytest exit $?" >> $S
for w in admm_kl_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2h_bench_$w.json 2> gpurun_out/r2h_bench_$w.err; echo "bench $w exit $?" >> $S
  python - gpurun_out/r2h_bench_$w.json <<'PY' >> $S 2>&1
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s', round(d['ms_per_step'], 3), 'ms', d['phase_ms'], d.get('objective_last'))
PY
done
K="regex:admm_|chol_inverse|gram_reduce|panel_gemm|residual_|sum_partials|fixed_planes|kl_data|ingest_data|row_planes3|zero_words"
CMD="python bench.py --workload ao_admm_65536x16384_k128 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 40 -c 300 --csv --log-file gpurun_out/r2h_launches_ao_admm.csv $CMD > gpurun_out/r2h_ncu_list1.log 2>&1
echo "ncu list ao exit $?" >> $S
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:admm_apply32_kernel" -s 12 -c 1 -o gpurun_out/full_apply32_r02h -f $CMD > gpurun_out/r2h_ncu_full1.log 2>&1
echo "ncu full apply32 exit $?" >> $S
ncu -i gpurun_out/full_apply32_r02h.ncu-rep --page raw --csv > gpurun_out/full_apply32_r02h.raw.csv 2>/dev/null
CMD="python bench.py --workload admm_kl_65536x16384_k128 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 70 -c 200 --csv --log-file gpurun_out/r2h_launches_admm_kl.csv $CMD > gpurun_out/r2h_ncu_list2.log 2>&1
echo "ncu list kl exit $?" >> $S
cat $S
