#!/bin/bash
# First gpurun call of round 2 (1 GPU, ~18 min): everything written after round 1's GPU budget was spent.
#   1. pytest -m gpu with per-test outcomes (-rA): which first_run_pending tests XPASS / XFAIL
#   2. experimental objective kernels (residual_n256.cu, residual_ts.cu): value + time against the default kernel
#      at the headline and C2 shapes (tools/selftest.py), then the headline bench with NMFB_RES_N256=1 / NMFB_RES_TS=1
#   3. fresh `ncu --set full` captures of the three kernels DESIGN.md section 9 asks about
#      (residual_kernel, residual_n256_kernel, residual_ts_kernel, kl_panel_kernel, kl_panel_ts_kernel): tensor-pipe activity against l1tex
#      throughput checks the shared-memory operand model of profiles/r01e_smem_operand_model.md
mkdir -p gpurun_out; S=gpurun_out/round2_first_summary.log; : > $S
timeout 1200 python -m pytest tests -m gpu -q -rA --timeout 600 --ignore=tests/test_gpu_zz_next_rows.py > gpurun_out/pytest_gpu_rA.log 2>&1; echo "pytest (validated files) exit $?" >> $S
tail -3 gpurun_out/pytest_gpu_rA.log >> $S
# every pending test in its own process (a faulting kernel poisons only its own CUDA context), --runxfail;
# also runs the headline bench with the e2e / time_to_tol legs (default objective kernel)
bash tools/gpu_pending.sh > gpurun_out/pending_stdout.log 2>&1
cat gpurun_out/pending_summary.log >> $S
for shape in "65536 65536 128" "32768 32768 64"; do
  timeout 600 python tools/selftest.py bench $shape obj > "gpurun_out/selftest_obj_${shape// /_}.log" 2>&1
  echo "selftest obj $shape exit $?" >> $S
  grep -E "objective" "gpurun_out/selftest_obj_${shape// /_}.log" >> $S
done
for var in NMFB_RES_N256 NMFB_RES_TS; do
  env $var=1 timeout 900 python bench.py --steps 30 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_headline_$var.json 2> gpurun_out/bench_headline_$var.err
  echo "bench headline $var=1 exit $?" >> $S
  python - <<PY >> $S
import json
for ln in open('gpurun_out/bench_headline_$var.json'):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s  phase_ms', d.get('phase_ms'), ' objective_last', d.get('objective_last'))
PY
done
for ts in 0 1; do   # ADMM / AO-ADMM objective reads two V planes (fixed-point ingest): residual_ts_kernel<2,2>
  for w in admm_65536x16384_k128 ao_admm_65536x16384_k128; do
    NMFB_RES_TS=$ts timeout 900 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_${w}_ts$ts.json 2> gpurun_out/bench_${w}_ts$ts.err
    echo "bench $w NMFB_RES_TS=$ts exit $?" >> $S
    python - <<PY >> $S
import json
for ln in open('gpurun_out/bench_${w}_ts$ts.json'):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s  phase_ms', d.get('phase_ms'), ' objective_last', d.get('objective_last'))
PY
  done
done
for ts in 0 1 2; do
  NMFB_KL_TS=$ts timeout 900 python bench.py --workload mur_kl_20000x50000_k128 --steps 30 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_kl_ts$ts.json 2> gpurun_out/bench_kl_ts$ts.err
  echo "bench mur_kl NMFB_KL_TS=$ts exit $?" >> $S
  python - <<PY >> $S
import json
for ln in open('gpurun_out/bench_kl_ts$ts.json'):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s  phase_ms', d.get('phase_ms'), ' objective_last', d.get('objective_last'))
PY
done
full() {  # env workload kernel-regex skip name
  CMD="python bench.py --workload $2 --steps 2 --warmup 3 --no-e2e --no-cpu"
  env $1 timeout 600 $CMD > gpurun_out/plain_$5.log 2>&1 &&
  env $1 timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$3" -s $4 -c 1 -o gpurun_out/full_$5 -f $CMD > gpurun_out/ncu_full_$5.log 2>&1
  echo "ncu full $5 exit $?" >> $S
  ncu -i gpurun_out/full_$5.ncu-rep --page raw --csv > gpurun_out/full_$5.raw.csv 2>/dev/null
}
full NMFB_RES_N256=0 mur_eu_65536x65536_k128 "residual_kernel" 3 residual_r02
full NMFB_RES_N256=1 mur_eu_65536x65536_k128 "residual_n256_kernel" 3 residual_n256_r02
full NMFB_RES_TS=1 mur_eu_65536x65536_k128 "residual_ts_kernel" 3 residual_ts_r02
full NMFB_KL_TS=0 mur_kl_20000x50000_k128 "kl_panel_kernel" 4 kl_panel_r02
full NMFB_KL_TS=1 mur_kl_20000x50000_k128 "kl_panel_ts_kernel" 4 kl_panel_ts_r02
full NMFB_KL_TS=2 mur_kl_20000x50000_k128 "kl_panel_ts2_kernel" 4 kl_panel_ts2_r02
python tools/ncu_summary.py > gpurun_out/ncu_summary_r02.md 2>/dev/null
cat $S
