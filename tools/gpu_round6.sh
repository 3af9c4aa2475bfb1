#!/bin/bash
mkdir -p gpurun_out
echo "== chunk huge, cluster 1"
NMFB_CLUSTER=1 NMFB_CHUNK_KB=100000 timeout 300 python tools/selftest.py bench 65536 65536 128 gemm 2>&1 | grep -E "split=0|rror"
echo "== chunk 8, cluster 1"
NMFB_CLUSTER=1 timeout 300 python tools/selftest.py bench 65536 65536 128 gemm 2>&1 | grep -E "split=0|rror"
NMFB_CLUSTER=1 timeout 300 python tools/selftest.py bench 65536 65536 128 gemm > gpurun_out/plain.log 2>&1 &&
NMFB_CLUSTER=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:panel_gemm -s 4 -c 1 -o gpurun_out/prof_pg2 -f python tools/selftest.py bench 65536 65536 128 gemm > gpurun_out/ncu_full.log 2>&1
echo "ncu exit $?"
