#!/bin/bash
# Round 2, call I (1 GPU): batched global loads in the k x k apply kernels (mur_eu_update, admm_apply, admm_apply32); kl_data branch-free.
mkdir -p gpurun_out; S=gpurun_out/r2i_summary.log; : > $S
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -k "admm or kl_branch or smoke or l2n or mur_eu or golden" > gpurun_out/r2i_pytest.log 2>&1; echo "pytest exit $?" >> $S
grep -E "passed|failed" gpurun_out/r2i_pytest.log | tail -3 >> $S
for w in mur_eu_65536x65536_k128 admm_65536x16384_k128 ao_admm_65536x16384_k128 admm_kl_65536x16384_k128 ao_admm_kl_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2i_bench_$w.json 2> gpurun_out/r2i_bench_$w.err; echo "bench $w exit $?" >> $S
  tail -2 gpurun_out/r2i_bench_$w.err >> $S
  python - gpurun_out/r2i_bench_$w.json <<'PY' >> $S 2>&1
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s', round(d['ms_per_step'], 3), 'ms', d['phase_ms'], d.get('objective_last'))
PY
done
K="regex:admm_|chol_inverse|gram_reduce|panel_gemm|residual_|sum_partials|fixed_planes|kl_data|ingest_data|row_planes3|zero_words|mur_eu_update"
CMD="python bench.py --workload ao_admm_65536x16384_k128 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 40 -c 300 --csv --log-file gpurun_out/r2i_launches_ao_admm.csv $CMD > gpurun_out/r2i_ncu_list1.log 2>&1
echo "ncu list ao exit $?" >> $S
CMD="python bench.py --workload admm_kl_65536x16384_k128 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 70 -c 200 --csv --log-file gpurun_out/r2i_launches_admm_kl.csv $CMD > gpurun_out/r2i_ncu_list2.log 2>&1
echo "ncu list kl exit $?" >> $S
CMD="python bench.py --workload mur_eu_65536x65536_k128 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 20 -c 200 --csv --log-file gpurun_out/r2i_launches_mur_eu.csv $CMD > gpurun_out/r2i_ncu_list3.log 2>&1
echo "ncu list mur exit $?" >> $S
cat $S
