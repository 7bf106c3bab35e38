#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
echo "== pytest multi"; timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 600 2>&1 | tail -4
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531"
echo "== sssp x$N 4096 check"; timeout 600 $TR tools/multi_sssp.py --width 4096 --height 4096 --check 2>gpurun_out/ssspc_x$N.err | grep '^{' | cut -c200-800; grep -i "error\|Traceback" gpurun_out/ssspc_x$N.err | head -3
for ce in 1 8 32; do
echo "== sssp x$N 8192 check_every $ce"; timeout 600 $TR tools/multi_sssp.py --width 8192 --height 8192 --check-every $ce 2>gpurun_out/sssp_x$N.err | grep '^{' | tee gpurun_out/sssp_x${N}_ce$ce.json | cut -c200-800; grep -i "error\|Traceback" gpurun_out/sssp_x$N.err | head -3
done
