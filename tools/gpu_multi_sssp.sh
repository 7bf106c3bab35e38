#!/bin/bash
# row-partitioned SSSP at N GPUs: parity tests, a checked 4096^2 run, the config-4 grid at a few windows
N=${1:-2}
mkdir -p gpurun_out
echo "== pytest multi"; timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 600 -k "sssp or empty" 2>&1 | tail -4
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531"
echo "== sssp x$N 4096 check"; timeout 600 $TR tools/multi_sssp.py --width 4096 --height 4096 --check 2>gpurun_out/ssspc_x$N.err | grep '^{' | cut -c200-800; grep -i "error\|Traceback" gpurun_out/ssspc_x$N.err | head -3
for w in ${2:-4 8 16}; do
echo "== sssp x$N 8192 window $w"; timeout 600 $TR tools/multi_sssp.py --width 8192 --height 8192 --window $w 2>gpurun_out/sssp_x$N.err | grep '^{' | tee gpurun_out/sssp_x${N}_w$w.json | cut -c200-800; grep -i "error\|Traceback" gpurun_out/sssp_x$N.err | head -3
done
