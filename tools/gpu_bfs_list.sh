#!/bin/bash
# launch list of the single-GPU BFS (RMAT 24): one ncu pass, after the plain run
mkdir -p gpurun_out
CMD="python tools/graph_profile.py 24 64 2"
timeout 300 $CMD > gpurun_out/graph_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/graph_plain.log; exit 1; }
grep bfs_scale gpurun_out/graph_plain.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'bfs_' -c 200 --csv --log-file gpurun_out/bfs_launches.csv $CMD > gpurun_out/ncu_bfs_list.log 2>&1
echo "launch list rc=$?"
