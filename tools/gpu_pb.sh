#!/bin/bash
# Blocked-schedule call: parity tests that include the blocked variant, then timing under a few geometries,
# the two roles alone (SHB_PB_SPLIT=p / c); then the new group / panel tests.
mkdir -p gpurun_out
echo "== pytest spmv"; timeout 240 python -m pytest tests/test_gpu_spmv.py -m gpu -x -q --timeout 60 > gpurun_out/pytest_spmv.txt 2>&1
rc=$?; tail -8 gpurun_out/pytest_spmv.txt
if [ $rc -ne 0 ]; then echo "parity tests failed or hung (rc=$rc): no timing"; exit 1; fi
echo "== default"; timeout 100 python tools/pb_bench.py --rs 0 2>&1 | cut -c1-400
for d in ${DEPTHS:-2}; do
echo "== depth $d"; SHB_PB_DEPTH=$d timeout 100 python tools/pb_bench.py --rs 0 --only-blocked 2>&1 | grep "blocked\|Error" | cut -c1-330
done
for w in ${MAXWS:-29000}; do
echo "== maxw $w"; SHB_PB_MAXW=$w timeout 100 python tools/pb_bench.py --rs 0 --only-blocked 2>&1 | grep "blocked\|Error" | cut -c1-330
done
echo "== produce only"; SHB_PB_SPLIT=p timeout 100 python tools/pb_bench.py --rs 0 --only-blocked --no-check 2>&1 | grep "blocked\|Error" | cut -c1-200
echo "== consume only"; SHB_PB_SPLIT=c timeout 100 python tools/pb_bench.py --rs 0 --only-blocked --no-check 2>&1 | grep "blocked\|Error" | cut -c1-200
echo "== pytest group + fullsize config2/5"; timeout 900 python -m pytest tests/test_gpu_group.py tests/test_gpu_fullsize.py -m gpu -x -q --timeout 400 -k "group or config2 or config5" 2>&1 | tail -8
echo "== group bench 1 gpu"; timeout 600 python tools/group_bench.py --gpus 1 --only pagerank 2>&1 | tail -3 | cut -c1-600
echo "== pagerank config5 no panels"; SHB_PANEL_MIN_BYTES=99999999999 timeout 600 python tools/group_bench.py --gpus 1 --only pagerank 2>&1 | tail -1 | cut -c1-300
