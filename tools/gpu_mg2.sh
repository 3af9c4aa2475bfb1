#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_g2.json 2> gpurun_out/bench_g2.err
echo "exit $?"; cut -c1-1800 gpurun_out/bench_g2.json; grep -v "^W0\|^\*\*\*\|OMP_NUM" gpurun_out/bench_g2.err | grep -B2 -A12 "Traceback" | head -60
