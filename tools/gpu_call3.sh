#!/bin/bash
mkdir -p gpurun_out
echo "== pytest multi"; timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 300 2>&1 | tail -15
echo "== sssp x1 1024 check"; timeout 300 python tools/multi_sssp.py --width 1024 --height 1024 --check 2>gpurun_out/s1.err | tee gpurun_out/sssp_x1_1k.json; tail -2 gpurun_out/s1.err
echo "== sssp x1 8192"; timeout 600 python tools/multi_sssp.py --width 8192 --height 8192 2>gpurun_out/s1.err | tee gpurun_out/sssp_x1_8k.json; tail -2 gpurun_out/s1.err
