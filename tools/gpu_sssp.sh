#!/bin/bash
# window sweep of the row-partitioned SSSP at N GPUs
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
echo "== check 2048 x$N"; timeout 300 $TR tools/multi_sssp.py --width 2048 --height 2048 --check 2>gpurun_out/sw.err | grep '^{' | cut -c200-700; grep -i "error\|Traceback" gpurun_out/sw.err | head -3
for w in 4 16 64 256; do
echo "== 8192 x$N window $w"; timeout 300 $TR tools/multi_sssp.py --width 8192 --height 8192 --window $w 2>gpurun_out/sw.err | grep '^{' | tee gpurun_out/sssp_x${N}_w$w.json | cut -c200-700; grep -i "error\|Traceback" gpurun_out/sw.err | head -3
done
