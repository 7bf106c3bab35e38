#!/bin/bash
# PageRank config 5, default (compressed-row) panels: plain run, then one ncu pass with per-launch counters
mkdir -p gpurun_out
timeout 200 python tools/group_bench.py --gpus 1 --only pagerank > gpurun_out/pr_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/pr_plain.log; exit 1; }
tail -1 gpurun_out/pr_plain.log | cut -c150-330
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors.sum,lts__t_sector_hit_rate.pct,l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum \
  --clock-control none -k regex:'spmv_|panel_' -s 30 -c 22 --csv --log-file gpurun_out/pr_panels_dcsr_ncu.csv \
  python tools/group_bench.py --gpus 1 --only pagerank > gpurun_out/pr_ncu.log 2>&1
echo "rc=$?"; wc -l gpurun_out/pr_panels_dcsr_ncu.csv
