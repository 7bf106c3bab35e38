#!/bin/bash
mkdir -p gpurun_out
echo "== pytest gpu multi"; timeout 1800 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 600 2>&1 | tail -15
echo "== bfs x1 s22 check"; timeout 600 python tools/multi_bfs.py --scale 22 --check 2>gpurun_out/bfs1c.err | tee gpurun_out/bfs_x1_s22.json; tail -3 gpurun_out/bfs1c.err
echo "== bfs x1 s24"; timeout 600 python tools/multi_bfs.py --scale 24 2>gpurun_out/bfs1.err | tee gpurun_out/bfs_x1_s24.json; tail -3 gpurun_out/bfs1.err
