#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.log
for t in gemm gram; do
  timeout 300 python tools/selftest.py $t > gpurun_out/st_$t.log 2>&1; echo "$t exit $?" >> gpurun_out/summary.log
  grep -c "OK " gpurun_out/st_$t.log; grep -E "BAD|Error|error|timeout" gpurun_out/st_$t.log | head -8
done
run() { echo "== $*"; env "$@" timeout 300 python tools/selftest.py bench 65536 65536 128 gemm 2>&1 | grep -E "split=0|H-side|rror|BAD"; }
run NMFB_CLUSTER=1
run NMFB_CLUSTER=2
run NMFB_CLUSTER=4
run NMFB_CLUSTER=1 NMFB_PREFETCH=8
run NMFB_CLUSTER=4 NMFB_PREFETCH=8
run NMFB_CLUSTER=1 NMFB_DAT_BOX=64
echo "== C2"
NMFB_CLUSTER=1 timeout 300 python tools/selftest.py bench 32768 32768 64 gemm 2>&1 | grep -E "split=0|H-side|rror"
NMFB_CLUSTER=4 timeout 300 python tools/selftest.py bench 32768 32768 64 gemm 2>&1 | grep -E "split=0|H-side|rror"
cat gpurun_out/summary.log
