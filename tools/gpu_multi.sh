#!/bin/bash
# N GPUs (gpurun --gpus N): sharded parity of every path, then sharded benches.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
timeout 900 $TR tools/mgpu_check.py > gpurun_out/mgpu_check_$N.log 2>&1; echo "mgpu_check exit $?"
grep -E "sharded|stop iteration|inner-iteration|MGPU|Error|error" gpurun_out/mgpu_check_$N.log | head -60
for w in mur_eu_65536x65536_k128 mur_kl_20000x50000_k128 admm_65536x16384_k128 ao_admm_65536x16384_k128; do
  timeout 600 $TR bench.py --gpus $N --workload $w --steps 20 --warmup 3 > gpurun_out/bench_g${N}_$w.json 2> gpurun_out/bench_g${N}_$w.err
  echo "bench $w x$N exit $?"
  python -c "
import json
d=json.load(open('gpurun_out/bench_g${N}_$w.json')); print('$w x$N', round(d['value'],1),'it/s', d['ms_per_step'], d['phase_ms'])" || grep -v "^W0\|^\*\*\*\|OMP_NUM" gpurun_out/bench_g${N}_$w.err | tail -12
done
