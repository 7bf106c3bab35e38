#!/bin/bash
# SHB_STREAM_LONG sweep: stream schedule on RMAT-22 / grid, and PageRank config 5 (column panels) per sweep
for v in 8 16 32 64; do
  echo "== SHB_STREAM_LONG=$v"
  SHB_STREAM_LONG=$v timeout 200 python tools/stream_bench.py 2>&1 | tail -1
  SHB_STREAM_LONG=$v timeout 300 python tools/group_bench.py --gpus 1 --only pagerank 2>&1 | tail -1 | python -c "
import sys, json
l = json.loads(sys.stdin.readline())
print({k: l[k] for k in ('ms_per_iteration', 'gteps', 'parity_ok', 'parity_rel_err')})"
done
