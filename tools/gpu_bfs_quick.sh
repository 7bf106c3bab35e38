#!/bin/bash
# BFS after a kernel change: small parity tests, the s24 timing (compacted pull at several grid sizes, thread per row),
# the full-size agreement test, the per-level trace
mkdir -p gpurun_out
echo "== pytest bfs (small)"; timeout 200 python -m pytest tests/test_gpu_multi.py tests/test_gpu_iterate.py -m gpu -x -q -k "bfs" 2>&1 | tail -3
echo "== timing (default)"; timeout 120 python tools/graph_profile.py 24 64 4 2>&1 | grep bfs_scale | tee gpurun_out/bfs_quick.log | tail -3
for m in 8 16 64 128; do echo "== timing (pull CTAs per SM $m)"; SHB_BFS_PULL_CTAS_PER_SM=$m timeout 120 python tools/graph_profile.py 24 64 3 2>&1 | grep bfs_scale | tail -2; done
echo "== timing (SHB_BFS_PULL=rows)"; SHB_BFS_PULL=rows timeout 120 python tools/graph_profile.py 24 64 3 2>&1 | grep bfs_scale | tail -2
echo "== pytest config3"; timeout 200 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "config3" 2>&1 | tail -3
echo "== trace"; timeout 200 python tools/multi_bfs.py --scale 24 --trace 2>/dev/null > gpurun_out/bfs_trace_x1.json; cut -c1-200 gpurun_out/bfs_trace_x1.json
