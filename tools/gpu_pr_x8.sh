#!/bin/bash
# PageRank config 5 on N GPUs: column panels off / 2 / 4 (default)
N=${N:-8}
echo "== no panels"; SHB_PANEL_MIN_BYTES=99999999999 timeout 200 python tools/group_bench.py --gpus $N --only pagerank 2>&1 | tail -1 | cut -c1-260
echo "== 2 panels"; SHB_PANEL_BYTES=134217728 timeout 200 python tools/group_bench.py --gpus $N --only pagerank 2>&1 | tail -1 | cut -c1-260
echo "== 4 panels (default)"; timeout 200 python tools/group_bench.py --gpus $N --only pagerank 2>&1 | tail -1 | cut -c1-260
