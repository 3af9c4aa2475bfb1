#!/bin/bash
# 8-GPU box: sharded parity at 8 ranks, headline scaling at 4 and 8, the other paths at 8.
mkdir -p gpurun_out
run() { # N workload steps
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29512"
  timeout 400 $TR bench.py --gpus $1 --workload $2 --steps $3 --warmup 3 > gpurun_out/bench_g$1_$2.json 2> gpurun_out/bench_g$1_$2.err
  echo "bench $2 x$1 exit $?"
  python -c "
import json
d=json.load(open('gpurun_out/bench_g$1_$2.json')); print('$2 x$1', round(d['value'],1),'it/s', round(d['ms_per_step'],3), d['phase_ms'])" || grep -v "^W0\|^\*\*\*\|OMP_NUM" gpurun_out/bench_g$1_$2.err | tail -12
}
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512"
timeout 420 $TR8 tools/mgpu_check.py > gpurun_out/mgpu_check_8.log 2>&1; echo "mgpu_check exit $?"
grep -E "sharded|stop iteration|inner-iteration|MGPU|Error|error" gpurun_out/mgpu_check_8.log | head -40
run 8 mur_eu_65536x65536_k128 30
run 4 mur_eu_65536x65536_k128 30
run 8 admm_65536x16384_k128 20
run 8 ao_admm_65536x16384_k128 20
run 8 mur_kl_20000x50000_k128 20
run 8 mur_eu_262144x131072_k256 10
