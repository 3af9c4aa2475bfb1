#!/bin/bash
for cfg in "8192 8192 128 2 0" "8192 8192 128 2 1" "32768 32768 128 2 1" "65536 65536 128 2 0" "65536 65536 64 2 0" "65536 65536 128 1 0"; do
  echo "== $cfg"; timeout 120 python tools/dbg_shard_obj.py $cfg 2>&1 | grep -E "objective|OK|rror" | cut -c1-200 | tail -5
done
echo "== sanitizer"
timeout 900 compute-sanitizer --tool memcheck --print-limit 5 python tools/dbg_shard_obj.py ${SAN:-8192 8192 128 2 1} 2>&1 | grep -v "^objective" | head -60 | cut -c1-250
