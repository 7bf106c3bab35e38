#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/grid_sweep.py 8192 2>&1 | tee gpurun_out/grid_sweep.jsonl | cut -c1-200
ncu --set full --clock-control none --import-source on -k regex:spmv_ -s 4 -c 2 -o gpurun_out/prof_grid python tools/grid_sweep.py 8192 auto > gpurun_out/ncu_grid.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/*.ncu-rep
