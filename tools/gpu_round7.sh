#!/bin/bash
run() { echo "== $*"; env "$@" NMFB_CLUSTER=1 timeout 300 python tools/selftest.py bench 65536 65536 128 gemm 2>&1 | grep -E "split=0|rror|BAD"; }
run NMFB_DAT_BOX=256
run NMFB_DAT_BOX=64
run NMFB_DAT_BOX=64 NMFB_PREFETCH=4
run NMFB_DAT_BOX=256 NMFB_PREFETCH=4
run NMFB_DAT_BOX=256 NMFB_PREFETCH=8
run NMFB_DAT_BOX=256 NMFB_PREFETCH=16
run NMFB_DAT_BOX=64 NMFB_PREFETCH=16
NMFB_PREFETCH=8 timeout 300 python tools/selftest.py gemm 2>&1 | grep -c "OK "
