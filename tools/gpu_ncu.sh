#!/bin/bash
# ncu evidence for bench.py, ONE ncu invocation per call (NCU=list | full | grid), always after the same command has
# exited 0 without ncu; numbers printed under ncu are never bench values.
#   list  launch list of the bench command (gpu__time_duration per launch)       -> gpurun_out/launches.csv
#   full  --set full capture of the dominant kernel of bench.py (config 2 SpMV)   -> gpurun_out/prof_spmv.ncu-rep
#   grid  --set full capture of the HBM-bound case (dense sweep, config-4 grid)   -> gpurun_out/prof_grid_stream.ncu-rep
mkdir -p gpurun_out
CMD="python bench.py --steps 5 --warmup 3 --extras 0 --cpu-steps 3"
case "${NCU:-list}" in
list)
  $CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/plain.log; exit 1; }
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
  echo "launch list rc=$?" ;;
full)
  $CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/plain.log; exit 1; }
  ncu --set full --clock-control none --import-source on -k regex:"spmv_vector_kernel|pb_fused_kernel" -s 3 -c 2 -o gpurun_out/prof_spmv $CMD > gpurun_out/ncu_full.log 2>&1
  echo "full capture rc=$?" ;;
grid)
  python tools/grid_sweep.py 8192 auto > gpurun_out/grid_plain.log 2>&1 || { echo "grid plain run failed"; tail -5 gpurun_out/grid_plain.log; exit 1; }
  ncu --set full --clock-control none --import-source on -k regex:spmv_stream_kernel -s 4 -c 2 -o gpurun_out/prof_grid_stream python tools/grid_sweep.py 8192 auto > gpurun_out/ncu_grid.log 2>&1
  echo "grid capture rc=$?" ;;
esac
ls -la gpurun_out | head -40
