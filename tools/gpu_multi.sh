#!/bin/bash
# Round 2, N-GPU call: peer-memory exchange -- sharded parity (tools/mgpu_check.py), sharded benches, one single-GPU line.
N=${1:-2}
mkdir -p gpurun_out; S=gpurun_out/r2c_summary_$N.log; : > $S
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
MGPU_CHECK_QUICK=${QUICK:-0} timeout 1500 $TR tools/mgpu_check.py > gpurun_out/r2c_mgpu_check_$N.log 2>&1; echo "mgpu_check x$N exit $?" >> $S
grep -E "sharded|stop iteration|inner-iteration|MGPU|Error|error" gpurun_out/r2c_mgpu_check_$N.log | head -80 >> $S
line() { python - "$1" <<'PY' >> "$2" 2>&1
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('  value', round(d['value'], 1), 'it/s', round(d['ms_per_step'], 3), 'ms', d['phase_ms'], d['config'].get('exchange', '')[:40], 'xchg', d.get('exchange_ms'), 'e2e', d.get('e2e'), 'ttt', d.get('time_to_tol'))
PY
}
for w in ${WORKLOADS:-mur_eu_65536x65536_k128 mur_kl_20000x50000_k128 admm_65536x16384_k128 ao_admm_65536x16384_k128}; do
  E2E="--no-e2e"; [ $w = mur_eu_65536x65536_k128 ] && [ "${NO_E2E:-0}" != 1 ] && E2E=""
  timeout ${BENCH_TIMEOUT:-600} $TR bench.py --gpus $N --workload $w --steps 20 --warmup 5 $E2E > gpurun_out/r2c_bench_g${N}_$w.json 2> gpurun_out/r2c_bench_g${N}_$w.err
  echo "bench $w x$N exit $?" >> $S
  line gpurun_out/r2c_bench_g${N}_$w.json $S || grep -v "^W0\|^\*\*\*\|OMP_NUM" gpurun_out/r2c_bench_g${N}_$w.err | tail -12 >> $S
done
if [ "${NOMC:-0}" = 1 ]; then
  NMFB_NO_MULTICAST=1 timeout ${BENCH_TIMEOUT:-600} $TR bench.py --gpus $N --workload mur_eu_65536x65536_k128 --steps 20 --warmup 5 --no-e2e > gpurun_out/r2c_bench_g${N}_nomc.json 2> gpurun_out/r2c_bench_g${N}_nomc.err
  echo "bench headline x$N without multicast exit $?" >> $S; line gpurun_out/r2c_bench_g${N}_nomc.json $S
fi
for w in ${SINGLE:-}; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2c_bench_g1_$w.json 2> gpurun_out/r2c_bench_g1_$w.err
  echo "bench $w x1 exit $?" >> $S; line gpurun_out/r2c_bench_g1_$w.json $S
done
if [ "${PYTEST_MULTI:-0}" = 1 ]; then
  timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r2c_pytest_multi_$N.log 2>&1; echo "pytest test_gpu_multi exit $?" >> $S
  tail -3 gpurun_out/r2c_pytest_multi_$N.log >> $S
fi
cat $S
