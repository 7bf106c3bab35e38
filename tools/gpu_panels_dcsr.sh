#!/bin/bash
# compressed-row panels: parity tests, then PageRank config 5 per sweep
timeout 600 python -m pytest tests/test_gpu_spmv.py tests/test_gpu_iterate.py tests/test_gpu_group.py tests/test_gpu_fullsize.py -m gpu -x -q --timeout 300 2>&1 | tail -6
for pb in 67108864 33554432; do
echo "== SHB_PANEL_BYTES=$pb"; SHB_PANEL_BYTES=$pb timeout 300 python tools/group_bench.py --gpus 1 --only pagerank 2>&1 | tail -1 | python -c "
import sys, json
l = json.loads(sys.stdin.readline())
print({k: l[k] for k in ('ms_per_iteration', 'gteps', 'parity_ok', 'parity_rel_err')})"
done
