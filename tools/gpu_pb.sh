#!/bin/bash
# Blocked-schedule call: parity tests that include the blocked variant, timing, per-kernel launch list.
mkdir -p gpurun_out
echo "== pytest spmv (split mode)"; SHB_PB_SPLIT=1 timeout 300 python -m pytest tests/test_gpu_spmv.py -m gpu -x -q --timeout 120 2>&1 | tail -8
echo "== pytest spmv (fused)"; timeout 300 python -m pytest tests/test_gpu_spmv.py -m gpu -x -q --timeout 120 2>&1 | tail -8
echo "== pb_bench split"; SHB_PB_SPLIT=1 timeout 300 python tools/pb_bench.py --rs 0 2>&1 | tail -6
echo "== pb_bench"; timeout 300 python tools/pb_bench.py --rs ${PB_RS:-0,1024,4096} 2>&1 | tail -20
for d in 2 8; do echo "== depth $d"; SHB_PB_DEPTH=$d timeout 300 python tools/pb_bench.py --rs 0 2>&1 | tail -2; done
echo "== ncu launch list"; SHB_PB_SPLIT=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:pb_fused --csv --log-file gpurun_out/pb_launches.csv python tools/pb_bench.py --rs 0 --reps 2 > gpurun_out/pb_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/pb_launches.csv")) if len(r) > 5]
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value")
for r in rows[1:]:
    print(r[ki][:50], r[vi])
PY
