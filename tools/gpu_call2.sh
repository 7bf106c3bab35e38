#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
echo "== pytest multi"; timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 600 2>&1 | tail -6
for ex in peer nccl; do
echo "== bfs x$N s22 check $ex"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 tools/multi_bfs.py --scale 22 --check --exchange $ex 2>gpurun_out/bfsc_${ex}_x$N.err | grep '^{' | tee gpurun_out/bfs_${ex}_x${N}_s22.json; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/bfsc_${ex}_x$N.err | tail -3
echo "== bfs x$N s24 $ex"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 tools/multi_bfs.py --scale 24 --exchange $ex 2>gpurun_out/bfs_${ex}_x$N.err | grep '^{' | tee gpurun_out/bfs_${ex}_x${N}_s24.json; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/bfs_${ex}_x$N.err | tail -3
done
