#!/bin/bash
# segmented-scan row phase of the stream schedule: parity, then A/B on RMAT-22 / grid / PageRank config 5
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_spmv.py tests/test_gpu_iterate.py -m gpu -x -q --timeout 120 2>&1 | tail -4
for v in 0 1; do
  echo "== SHB_STREAM_SCAN=$v"
  SHB_STREAM_SCAN=$v timeout 200 python tools/stream_bench.py 2>&1 | tail -1
  SHB_STREAM_SCAN=$v timeout 300 python tools/group_bench.py --gpus 1 --only pagerank 2>&1 | tail -1 | python -c "
import sys, json
l = json.loads(sys.stdin.readline())
print({k: l[k] for k in ('ms_per_iteration', 'gteps', 'parity_ok', 'parity_rel_err')})"
done
