#!/bin/bash
# ncu evidence for the graph drivers on one GPU: launch list of BFS (RMAT 24) + SSSP (8192^2 grid), one full capture
# of the BFS pull kernel (the two big levels) and one of the persistent near-far SSSP kernel.  Plain run first.
mkdir -p gpurun_out
CMD="python tools/graph_profile.py 24 8192 2"
timeout 300 $CMD > gpurun_out/graph_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/graph_plain.log; exit 1; }
cat gpurun_out/graph_plain.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'bfs_|sssp_' -c 400 --csv --log-file gpurun_out/graph_launches.csv $CMD > gpurun_out/ncu_graph_list.log 2>&1
echo "launch list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:bfs_block_pull_kernel -c 3 -o gpurun_out/prof_bfs_pull python tools/graph_profile.py 24 64 1 > gpurun_out/ncu_bfs_full.log 2>&1
echo "bfs capture rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:sssp_nearfar_kernel -c 1 -o gpurun_out/prof_sssp python tools/graph_profile.py 10 8192 1 > gpurun_out/ncu_sssp_full.log 2>&1
echo "sssp capture rc=$?"; tail -3 gpurun_out/ncu_sssp_full.log
ls -la gpurun_out | grep -E "prof_|graph_"
