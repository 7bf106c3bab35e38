#!/bin/bash
# PageRank config 5 per sweep by x-slice size of the (compressed-row) column panels, N GPUs
N=${N:-1}
for pb in ${SIZES:-8388608 16777216 25165824 33554432}; do
echo "== SHB_PANEL_BYTES=$pb"; SHB_GROUP_PANEL_BYTES=$pb SHB_PANEL_BYTES=$pb timeout 300 python tools/group_bench.py --gpus $N --only pagerank 2>&1 | tail -1 | python -c "
import sys, json
l = json.loads(sys.stdin.readline())
print({k: l[k] for k in ('ms_per_iteration', 'gteps', 'parity_ok', 'parity_rel_err')})"
done
