#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
echo "== KL tests (1 GPU)"; timeout 600 python -m pytest tests/test_gpu_kl.py -m gpu -q --timeout 120 2>&1 | tail -15
timeout 300 python bench.py --workload mur_kl_20000x50000_k128 --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_kl.json 2> gpurun_out/bench_kl.err; echo "bench kl exit $?"; cut -c1-900 gpurun_out/bench_kl.json; tail -3 gpurun_out/bench_kl.err
for w in mur_eu_4096x4096_k64 mur_eu_32768x32768_k64; do
  timeout 300 $TR bench.py --gpus 2 --workload $w --steps 5 --warmup 3 > gpurun_out/dbg_$w.json 2> gpurun_out/dbg_$w.err; echo "x2 $w exit $?"; cut -c1-300 gpurun_out/dbg_$w.json
done
CUDA_LAUNCH_BLOCKING=1 timeout 300 $TR bench.py --gpus 2 --workload mur_eu_65536x65536_k128 --steps 2 --warmup 3 > gpurun_out/dbg_blocking.json 2> gpurun_out/dbg_blocking.err; echo "blocking exit $?"
grep -E "^\[rank[01]\]:   File|^\[rank[01]\]: .*Error|nmf_b200 error" gpurun_out/dbg_blocking.err | head -30
