#!/bin/bash
# PageRank on BASELINE config 5 (RMAT scale 26, 2^30 non-zeros) on one GPU: ms per sweep by x-slice size of the column
# panels, then one ncu pass (durations + DRAM / L2 / request counters per launch) of the default setting.
mkdir -p gpurun_out
for pb in 16777216 33554432 67108864 134217728 1073741824; do
  echo "== SHB_PANEL_BYTES=$pb"
  SHB_PANEL_BYTES=$pb timeout 300 python tools/group_bench.py --gpus 1 --only pagerank 2>&1 | tail -1 | python -c "
import sys, json
l = json.loads(sys.stdin.readline())
print({k: l[k] for k in ('ms_per_iteration', 'gteps', 'parity_ok', 'parity_rel_err')})"
done
PB=${NCU_PANEL_BYTES:-33554432}
echo "== ncu (SHB_PANEL_BYTES=$PB)"
SHB_PANEL_BYTES=$PB timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors.sum,lts__t_sector_hit_rate.pct,l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum \
  --clock-control none -k regex:spmv_ -s 40 -c 40 --csv --log-file gpurun_out/pr_panels_ncu.csv \
  python tools/group_bench.py --gpus 1 --only pagerank > gpurun_out/pr_panels_ncu.log 2>&1
echo "rc=$?"; wc -l gpurun_out/pr_panels_ncu.csv
