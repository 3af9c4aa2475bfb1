#!/bin/bash
# Round 2, 1-GPU call: the no-allocation Gram entry point alone in a fresh process, then the whole GPU suite (it is
# collected last there), then the default bench line.
mkdir -p gpurun_out; S=gpurun_out/r2o_summary.log; : > $S
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -s > gpurun_out/r2o_no_alloc_alone.log 2>&1; echo "no_alloc alone exit $?" >> $S
grep -E "static vs|passed|failed|Error" gpurun_out/r2o_no_alloc_alone.log >> $S
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2o_gpu_suite.log 2>&1; echo "gpu suite exit $?" >> $S
tail -5 gpurun_out/r2o_gpu_suite.log >> $S
timeout 900 python bench.py > gpurun_out/r2o_bench_default.json 2> gpurun_out/r2o_bench_default.err; echo "bench exit $?" >> $S
cut -c1-1500 gpurun_out/r2o_bench_default.json >> $S
cat $S
