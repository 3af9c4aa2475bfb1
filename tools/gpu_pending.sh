#!/bin/bash
# First GPU run of everything written without a GPU at hand (tests/test_gpu_zz_next_rows.py: l2n prox,
# KL branch of admm / ao_admm, full-size property tests, edge cases): every test in its OWN process (a
# faulting kernel poisons its CUDA context, not the others'), with --runxfail so that failures show their
# traceback.  Then the bench legs added late (time_to_tol; the N > 1 e2e leg needs gpu_multi.sh).
mkdir -p gpurun_out; S=gpurun_out/pending_summary.log; : > $S
IDS=$(python -m pytest tests/test_gpu_zz_next_rows.py -m gpu --collect-only -q 2>/dev/null | grep "::")
for id in $IDS; do
  name=$(echo "$id" | sed 's/[^A-Za-z0-9_]/_/g' | cut -c1-120)
  timeout 900 python -m pytest "$id" -m gpu --runxfail -q --timeout 600 > gpurun_out/pending_$name.log 2>&1
  rc=$?
  echo "$rc $id" >> $S
  if [ $rc -ne 0 ]; then tail -40 gpurun_out/pending_$name.log; fi
done
timeout 900 python bench.py --steps 30 --warmup 3 --no-cpu > gpurun_out/bench_headline_ttt.json 2> gpurun_out/bench_headline_ttt.err
echo "bench exit $?" >> $S
python - <<'PY'
import json
for ln in open('gpurun_out/bench_headline_ttt.json'):
    if ln.startswith('{'):
        d = json.loads(ln)
        print('value', round(d['value'], 1), 'e2e', d.get('e2e'), 'time_to_tol', d.get('time_to_tol'))
PY
cat $S
