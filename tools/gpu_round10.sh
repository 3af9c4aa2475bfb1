#!/bin/bash
for cs in 1 2 4; do echo "== res cluster $cs"; NMFB_RES_CLUSTER=$cs timeout 300 python tools/selftest.py bench 65536 65536 128 obj 2>&1 | grep -E "objective|rror"; done
timeout 300 python tools/selftest.py objective 2>&1 | tail -2
