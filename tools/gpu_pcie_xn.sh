#!/bin/bash
# Host<->device copy ceiling with N GPUs busy at once (one process per GPU, 64 MB pinned buffers both ways):
# what the e2e leg of bench.py --gpus N can reach on this box.
N=${N:-8}
mkdir -p gpurun_out
for i in $(seq 0 $((N-1))); do
  CUDA_VISIBLE_DEVICES=$i timeout 120 python tools/pcie_bench.py --reps 40 > gpurun_out/pcie_x${N}_$i.json 2>/dev/null &
done
wait
python - <<PY
import json
n=$N
rows=[json.load(open(f"gpurun_out/pcie_x{n}_{i}.json")) for i in range(n)]
out={"n_gpus": n, "note": "all processes run the four phases at about the same time (not barrier-aligned)"}
for k in ("h2d_alone","d2h_alone","both","both_with_device_traffic"):
    out[k]={"ms_per_step_max": max(r[k]["ms_per_step"] for r in rows), "ms_per_step_min": min(r[k]["ms_per_step"] for r in rows),
            "aggregate_gbs_per_direction": sum(r[k]["gbs_per_direction"] for r in rows)}
print(json.dumps(out))
open(f"gpurun_out/pcie_x{n}.json","w").write(json.dumps(out))
PY
