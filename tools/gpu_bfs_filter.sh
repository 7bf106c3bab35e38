#!/bin/bash
# BFS after the coarse frontier filter: parity (small + full size), timing at scale 24, per-level trace
mkdir -p gpurun_out
echo "== pytest bfs"; timeout 400 python -m pytest tests/test_gpu_multi.py tests/test_gpu_iterate.py tests/test_gpu_group.py tests/test_gpu_host.py -m gpu -x -q -k "bfs or group" --timeout 200 2>&1 | tail -3
echo "== timing"; timeout 120 python tools/graph_profile.py 24 64 4 2>&1 | grep bfs_scale | tail -3
echo "== pytest config3"; timeout 300 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "config3" 2>&1 | tail -3
echo "== trace"; timeout 200 python tools/multi_bfs.py --scale 24 --trace 2>/dev/null > gpurun_out/bfs_trace_x1.json; python -c "
import json; d=json.load(open('gpurun_out/bfs_trace_x1.json')); print(d['ms'], d['gteps']); print(d['trace'])"
