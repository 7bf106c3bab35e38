#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
echo "== pytest multi"; timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 600 2>&1 | tail -6
echo "== sssp x$N 2048 check"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 tools/multi_sssp.py --width 2048 --height 2048 --check 2>gpurun_out/ssspc_x$N.err | grep '^{' | tee gpurun_out/sssp_x${N}_2k.json; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/ssspc_x$N.err | tail -3
echo "== sssp x$N 8192"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 tools/multi_sssp.py --width 8192 --height 8192 2>gpurun_out/sssp_x$N.err | grep '^{' | tee gpurun_out/sssp_x${N}_8k.json; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/sssp_x$N.err | tail -3
echo "== bfs x$N s24 peer"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 tools/multi_bfs.py --scale 24 --trace 2>gpurun_out/bfs_x$N.err | grep '^{' | tee gpurun_out/bfs_x${N}_s24.json; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/bfs_x$N.err | tail -3
