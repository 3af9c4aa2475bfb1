#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.log
for t in gemm gram; do
  timeout 300 python tools/selftest.py $t > gpurun_out/st_$t.log 2>&1; echo "$t exit $?" >> gpurun_out/summary.log
  grep -c "OK " gpurun_out/st_$t.log; grep -E "BAD|Error|error|timeout" gpurun_out/st_$t.log | head -8
done
for cs in 1 2 4; do
  echo "== cluster $cs"
  NMFB_CLUSTER=$cs timeout 300 python tools/selftest.py bench 65536 65536 128 gemm 2>&1 | grep -E "split=0|H-side|rror"
done
echo "== chunk 4 / 16, cluster 4"
NMFB_CHUNK_KB=4 timeout 300 python tools/selftest.py bench 65536 65536 128 gemm 2>&1 | grep -E "split=0|rror"
NMFB_CHUNK_KB=16 timeout 300 python tools/selftest.py bench 65536 65536 128 gemm 2>&1 | grep -E "split=0|rror"
echo "== C2"
timeout 300 python tools/selftest.py bench 32768 32768 64 gemm 2>&1 | grep -E "split=0|H-side|rror"
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/summary.log
tail -8 gpurun_out/pytest_gpu.log
cat gpurun_out/summary.log
