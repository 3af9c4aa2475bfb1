#!/bin/bash
for rep in 1 2 3; do
for cfg in "8192 8192 128 2 0" "8192 8192 128 2 1" "65536 65536 128 1 0" "16384 16384 64 2 1" "8192 8192 256 2 0"; do
  echo "== $cfg"; timeout 120 python tools/dbg_shard_obj.py $cfg 2>&1 | grep -E "objective 2|OK|rror" | cut -c1-160 | tail -3
done
done
