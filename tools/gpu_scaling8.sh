#!/bin/bash
# Round 2, 8-GPU call: sharded parity at 8 ranks (frozen oracle runs), headline with e2e / time_to_tol, config 5, other solvers.
N=${1:-8}
mkdir -p gpurun_out; S=gpurun_out/s8_summary_$N.log; : > $S
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
line() { python - "$1" <<'PY' >> "$2" 2>&1
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        r = d.get('roofline') or {}
        print('  value', round(d['value'], 2), 'it/s', round(d['ms_per_step'], 3), 'ms', d['phase_ms'], 'roof', r.get('bound'), round(r.get('frac') or 0, 3),
              d['config'].get('exchange', '')[:24], 'xchg', d.get('exchange_ms'), 'e2e', d.get('e2e'), 'ttt', d.get('time_to_tol'))
PY
}
for w in ${WORKLOADS:-mur_eu_65536x65536_k128 mur_eu_262144x131072_k256 admm_65536x16384_k128 ao_admm_65536x16384_k128 mur_kl_20000x50000_k128}; do
  E2E="--no-e2e"; [ $w = mur_eu_65536x65536_k128 ] && [ "${NO_E2E:-0}" != 1 ] && E2E=""
  timeout ${BENCH_TIMEOUT:-600} $TR bench.py --gpus $N --workload $w --steps 20 --warmup 5 $E2E > gpurun_out/s8_bench_g${N}_$w.json 2> gpurun_out/s8_bench_g${N}_$w.err
  echo "bench $w x$N exit $?" >> $S
  line gpurun_out/s8_bench_g${N}_$w.json $S || grep -v "^W0\|^\*\*\*\|OMP_NUM" gpurun_out/s8_bench_g${N}_$w.err | tail -12 >> $S
done
[ "${SKIP_CHECK:-0}" = 1 ] && { cat $S; exit 0; }
MGPU_CHECK_QUICK=1 timeout 900 $TR tools/mgpu_check.py > gpurun_out/s8_mgpu_check_$N.log 2>&1; echo "mgpu_check x$N exit $?" >> $S
grep -E "sharded|stop iteration|inner-iteration|MGPU|Error|error" gpurun_out/s8_mgpu_check_$N.log | head -60 >> $S
cat $S
