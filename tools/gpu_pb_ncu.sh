#!/bin/bash
mkdir -p gpurun_out
SHB_PB_SPLIT=1 SHB_PB_MAXW=${MAXW:-30000} timeout 900 ncu --set full --clock-control none --import-source on -k regex:pb_fused -s 2 -c 2 -f -o gpurun_out/pb_split python tools/pb_bench.py --rs 0 --reps 2 > gpurun_out/pb_ncu_full.log 2>&1
echo rc=$?
ls -la gpurun_out/*.ncu-rep
