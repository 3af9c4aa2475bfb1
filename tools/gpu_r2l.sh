#!/bin/bash
mkdir -p gpurun_out; S=gpurun_out/r2l_summary.log; : > $S
timeout 600 python tools/diag_admm_step.py 8 > gpurun_out/r2l_diag_step.log 2>&1; echo "diag step exit $?" >> $S; cat gpurun_out/r2l_diag_step.log >> $S
timeout 900 python tools/diag_parity.py > gpurun_out/r2l_diag.log 2>&1; echo "diag exit $?" >> $S
grep -E "max relObj|after" gpurun_out/r2l_diag.log | cut -c1-330 >> $S
cat $S
