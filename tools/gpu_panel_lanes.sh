#!/bin/bash
# PageRank config 5: the panels under the vector schedule with L lanes per row, against the stream schedule (default)
for L in 1 2 4; do
  echo "== SHB_PANEL_LANES=$L"
  SHB_PANEL_LANES=$L timeout 300 python tools/group_bench.py --gpus 1 --only pagerank 2>&1 | tail -1 | python -c "
import sys, json
l = json.loads(sys.stdin.readline())
print({k: l[k] for k in ('ms_per_iteration', 'gteps', 'parity_ok', 'parity_rel_err')})"
done
