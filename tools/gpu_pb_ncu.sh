#!/bin/bash
# Blocked schedule: parity, timing, then ncu --set full captures of the fused launch and of the two roles alone
# (SHB_PB_SPLIT=1 runs producer and consumer as two launches over a full-size ring).
mkdir -p gpurun_out
echo "== pytest spmv"; timeout 240 python -m pytest tests/test_gpu_spmv.py -m gpu -x -q --timeout 60 > gpurun_out/pytest_spmv.txt 2>&1
rc=$?; tail -5 gpurun_out/pytest_spmv.txt
if [ $rc -ne 0 ]; then echo "parity tests failed or hung (rc=$rc): no timing"; exit 1; fi
echo "== default"; timeout 100 python tools/pb_bench.py --rs 0 2>&1 | cut -c1-420
echo "== produce only"; SHB_PB_SPLIT=p timeout 100 python tools/pb_bench.py --rs 0 --only-blocked --no-check 2>&1 | grep "blocked\|Error" | cut -c1-120
echo "== consume only"; SHB_PB_SPLIT=c timeout 100 python tools/pb_bench.py --rs 0 --only-blocked --no-check 2>&1 | grep "blocked\|Error" | cut -c1-120
CMD="python tools/pb_bench.py --rs 0 --only-blocked --no-check --reps 1"
# one ncu capture per call: NCU=fused (default) | split | none
case "${NCU:-fused}" in
fused)
  echo "== ncu fused"
  timeout 100 $CMD > gpurun_out/pb_plain.log 2>&1 && \
  timeout 500 ncu --set full --clock-control none --import-source on -k regex:pb_fused -s 2 -c 1 -o gpurun_out/prof_pb_fused $CMD > gpurun_out/ncu_pb_fused.log 2>&1
  echo "rc=$?" ;;
split)
  echo "== ncu split"
  export SHB_PB_SPLIT=1
  timeout 100 $CMD > gpurun_out/pb_plain_split.log 2>&1 && \
  timeout 500 ncu --set full --clock-control none --import-source on -k regex:pb_fused -s 4 -c 2 -o gpurun_out/prof_pb_split $CMD > gpurun_out/ncu_pb_split.log 2>&1
  echo "rc=$?" ;;
esac
ls -la gpurun_out | grep prof_pb
