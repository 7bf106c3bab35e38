#!/bin/bash
# BFS after a kernel change: small parity tests, the s24 timing, then the full-size agreement test
mkdir -p gpurun_out
echo "== pytest bfs (small)"; timeout 200 python -m pytest tests/test_gpu_multi.py tests/test_gpu_iterate.py -m gpu -x -q -k "bfs" 2>&1 | tail -3
echo "== timing"; timeout 120 python tools/graph_profile.py 24 64 4 2>&1 | tee gpurun_out/bfs_quick.log | tail -6
echo "== pytest config3"; timeout 200 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "config3" 2>&1 | tail -3
