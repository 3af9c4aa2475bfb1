#!/bin/bash
# Round 2, call A (1 GPU): A/B of the objective-kernel variants and the MUR-KL TS variants, fresh ncu --set full.
mkdir -p gpurun_out; S=gpurun_out/r2a_summary.log; : > $S
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv >> $S
for shape in "65536 65536 128" "32768 32768 64"; do
  timeout 600 python tools/selftest.py bench $shape obj > "gpurun_out/selftest_obj_${shape// /_}.log" 2>&1
  echo "selftest obj $shape exit $?" >> $S
  grep -E "objective" "gpurun_out/selftest_obj_${shape// /_}.log" >> $S
done
line() { python - "$1" <<'PY' >> gpurun_out/r2a_summary.log
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s  phase_ms', d.get('phase_ms'), ' objective_last', d.get('objective_last'))
PY
}
for var in NMFB_NONE NMFB_RES_N256 NMFB_RES_TS; do
  env $var=1 timeout 900 python bench.py --steps 30 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_headline_$var.json 2> gpurun_out/bench_headline_$var.err
  echo "bench headline $var=1 exit $?" >> $S; line gpurun_out/bench_headline_$var.json
done
for ts in 0 1; do
  for w in admm_65536x16384_k128 ao_admm_65536x16384_k128; do
    NMFB_RES_TS=$ts timeout 900 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_${w}_ts$ts.json 2> gpurun_out/bench_${w}_ts$ts.err
    echo "bench $w NMFB_RES_TS=$ts exit $?" >> $S; line gpurun_out/bench_${w}_ts$ts.json
  done
done
for ts in 0 1 2; do
  NMFB_KL_TS=$ts timeout 900 python bench.py --workload mur_kl_20000x50000_k128 --steps 30 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_kl_ts$ts.json 2> gpurun_out/bench_kl_ts$ts.err
  echo "bench mur_kl NMFB_KL_TS=$ts exit $?" >> $S; line gpurun_out/bench_kl_ts$ts.json
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
full NMFB_NONE=1 admm_65536x16384_k128 "admm_apply_kernel" 5 admm_apply_r02
full NMFB_NONE=1 admm_65536x16384_k128 "chol_inverse_kernel" 3 chol_r02
python tools/ncu_summary.py > gpurun_out/ncu_summary_r02.md 2>/dev/null
cat $S
