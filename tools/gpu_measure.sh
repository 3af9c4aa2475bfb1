#!/bin/bash
# 1-GPU measurement pass: GPU suite, smoke(), the bench lines of every workload (default line with e2e / time_to_tol /
# CPU arm), the reference arm, the ncu launch list of the default command and one `ncu --set full` capture per shipped
# hot kernel (each only after the same command has exited 0 without ncu).  Everything lands in gpurun_out/.
mkdir -p gpurun_out; S=gpurun_out/measure_summary.log; : > $S
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv,noheader >> $S
if [ "${SKIP_SUITE:-0}" != 1 ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/measure_gpu_suite.log 2>&1; echo "gpu suite exit $?" >> $S
  tail -3 gpurun_out/measure_gpu_suite.log >> $S
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/measure_smoke.log 2>&1; echo "smoke exit $?" >> $S
  tail -2 gpurun_out/measure_smoke.log >> $S
fi
line() { python - "$1" <<'PY' >> gpurun_out/measure_summary.log 2>&1
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        r = d.get('roofline') or {}
        e = d.get('e2e') or {}
        print('  value', round(d.get('value') or 0, 2), d.get('unit'), 'ms', round(d.get('ms_per_step') or 0, 3), d.get('phase_ms'),
              '| roof', r.get('kernel'), r.get('bound'), round(r.get('frac') or 0, 3), '| e2e', e.get('value'),
              '| ttt', (d.get('time_to_tol') or {}).get('seconds'), (d.get('time_to_tol') or {}).get('iterations'),
              '| cpu', d.get('cpu_baseline'), '| clocks', d.get('clocks'))
PY
}
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default exit $?" >> $S
line gpurun_out/bench_default.json
if [ "${REFERENCE_ARM:-0}" = 1 ]; then
  timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_arm.json 2> gpurun_out/bench_reference_arm.err
  echo "bench --impl reference exit $?" >> $S; cut -c1-600 gpurun_out/bench_reference_arm.json >> $S
fi
for w in ${WORKLOADS:-mur_eu_32768x32768_k64 mur_kl_20000x50000_k128 admm_65536x16384_k128 ao_admm_65536x16384_k128}; do
  timeout 900 python bench.py --workload $w --steps 20 --warmup 3 ${CPU_FLAG:---no-cpu} > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
  echo "bench $w exit $?" >> $S; line gpurun_out/bench_$w.json
done
for w in ${WORKLOADS_NOCPU:-mur_eu_65536x32768_k256 admm_kl_65536x16384_k128 ao_admm_kl_65536x16384_k128}; do
  timeout 900 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
  echo "bench $w exit $?" >> $S; line gpurun_out/bench_$w.json
done
# launch list of the default command (cold-cache, serialised per-launch times: the SHARES are what is compared)
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 600 $CMD > gpurun_out/plain_default.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_default.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launch list exit $?" >> $S
full() {  # workload kernel-regex skip name
  CMD="python bench.py --workload $1 --steps 2 --warmup 3 --no-e2e --no-cpu"
  timeout 600 $CMD > gpurun_out/plain_$4.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$2" -s $3 -c 1 -o gpurun_out/full_$4 -f $CMD > gpurun_out/ncu_full_$4.log 2>&1
  echo "ncu full $4 exit $?" >> $S
  ncu -i gpurun_out/full_$4.ncu-rep --page raw --csv > gpurun_out/full_$4.raw.csv 2>/dev/null
  rm -f gpurun_out/full_$4.ncu-rep
}
if [ "${SKIP_NCU:-0}" != 1 ]; then
  full mur_eu_65536x65536_k128 "panel_gemm_kernel" 4 panel_gemm
  full mur_eu_65536x65536_k128 "residual_n256_kernel" 3 residual_n256
  full mur_kl_20000x50000_k128 "kl_panel_ts2_kernel" 4 kl_panel_ts2
  full admm_65536x16384_k128 "admm_apply_kernel" 5 admm_apply
  full admm_kl_65536x16384_k128 "kl_data_kernel" 2 kl_data
  python tools/ncu_summary.py full gpurun_out/full_*.raw.csv > gpurun_out/ncu_full_summary.md 2>> $S
  python tools/ncu_summary.py launches gpurun_out/launches_default.csv > gpurun_out/ncu_launches_summary.md 2>> $S
fi
cat $S
