#!/bin/bash
# Full measurement pass on the current kernels: GPU tests, all bench workloads, ncu launch lists,
# ncu --set full captures of each dominant kernel (raw CSV pages written next to the reports).
mkdir -p gpurun_out; rm -f gpurun_out/summary.log
S=gpurun_out/summary.log
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $S
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 30 --warmup 3 > gpurun_out/bench_headline.json 2> gpurun_out/bench_headline.err; echo "bench exit $?" >> $S
cut -c1-2500 gpurun_out/bench_headline.json; tail -3 gpurun_out/bench_headline.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench_ref exit $?" >> $S
for w in mur_eu_32768x32768_k64 mur_kl_20000x50000_k128 admm_65536x16384_k128 ao_admm_65536x16384_k128; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w exit $?" >> $S
  python -c "
import json
d=json.load(open('gpurun_out/bench_$w.json')); print('$w', round(d['value'],1),'it/s e2e', d['e2e'] and round(d['e2e']['value'],1), d['phase_ms'], d['roofline'].get('frac'))" || tail -5 gpurun_out/bench_$w.err
done
# ---- ncu: launch lists (only after the plain command exited 0) ----
for w in mur_eu_65536x65536_k128 mur_kl_20000x50000_k128 admm_65536x16384_k128 ao_admm_65536x16384_k128; do
  CMD="python bench.py --workload $w --steps 2 --warmup 3 --no-e2e --no-cpu"
  timeout 600 $CMD > gpurun_out/plain_$w.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:panel_gemm|residual_kernel|_update_kernel|gram_reduce|sum_partials|kl_panel|rowsum|admm_|chol_inverse|fixed_planes|objective_simt" -c 600 --csv --log-file gpurun_out/launches_$w.csv $CMD > gpurun_out/ncu_list_$w.log 2>&1
  echo "ncu list $w exit $?" >> $S
done
# ---- ncu --set full: one launch of each dominant kernel ----
full() {  # workload kernel-regex skip name
  CMD="python bench.py --workload $1 --steps 2 --warmup 3 --no-e2e --no-cpu"
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$2" -s $3 -c 1 -o gpurun_out/full_$4 -f $CMD > gpurun_out/ncu_full_$4.log 2>&1
  echo "ncu full $4 exit $?" >> $S
  ncu -i gpurun_out/full_$4.ncu-rep --page raw --csv > gpurun_out/full_$4.raw.csv 2>/dev/null
}
full mur_eu_65536x65536_k128 panel_gemm_kernel 9 panel_gemm
full mur_eu_65536x65536_k128 residual_kernel 3 residual
full mur_eu_65536x65536_k128 mur_eu_update 4 mur_eu_update
full mur_kl_20000x50000_k128 kl_panel_kernel 4 kl_panel
full admm_65536x16384_k128 admm_apply_kernel 5 admm_apply
full admm_65536x16384_k128 panel_gemm_exact 5 panel_gemm_exact
cat $S
