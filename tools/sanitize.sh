#!/bin/bash
# compute-sanitizer over every solver and branch at small shapes (tools/sanitize_target.py).  One log per tool under
# gpurun_out/; the summary lists exit codes and the sanitizer's own error counts.  Usage: bash tools/sanitize.sh [tools...]
mkdir -p gpurun_out; S=gpurun_out/sanitize_summary.log; : > $S
for tool in ${@:-memcheck racecheck synccheck}; do
  timeout ${SANITIZE_TIMEOUT:-600} compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 \
      python tools/sanitize_target.py > gpurun_out/sanitize_$tool.log 2>&1
  echo "compute-sanitizer $tool exit $? (9 = errors reported, 124 = timeout)" >> $S
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|^[0-9]+x[0-9]+ k=" gpurun_out/sanitize_$tool.log | tail -20 >> $S
done
cat $S
