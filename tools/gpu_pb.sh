#!/bin/bash
# Blocked-schedule call: parity tests that include the blocked variant, timing under a few geometries.
mkdir -p gpurun_out
echo "== pytest spmv (fused)"; timeout 300 python -m pytest tests/test_gpu_spmv.py -m gpu -x -q --timeout 120 2>&1 | tail -8
for w in ${MAXWS:-40960 30000 19000}; do
echo "== maxw $w"; SHB_PB_MAXW=$w timeout 300 python tools/pb_bench.py --rs 0 2>&1 | grep "blocked\|Error" | cut -c1-330
done
echo "== split"; SHB_PB_SPLIT=1 SHB_PB_MAXW=${SPLITW:-30000} timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:pb_fused --csv --log-file gpurun_out/pb_launches.csv python tools/pb_bench.py --rs 0 --reps 2 > gpurun_out/pb_ncu.log 2>&1
python - <<'PY'
import csv
rows = [r for r in csv.reader(open("gpurun_out/pb_launches.csv")) if len(r) > 5]
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value")
for r in rows[1:5]:
    print(r[ki][:50], r[vi])
PY
