#!/bin/bash
# Multi-GPU gpurun call (N GPUs of one box): the >=2-GPU tests, bench.py under torchrun, row-partitioned PageRank.
N=${1:-2}; SCALE=${2:-24}; CHECK_SCALE=${3:-22}
mkdir -p gpurun_out
nvidia-smi -L | head -8
echo "== pytest multi"; timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 600 2>&1 | tail -4
echo "== bench x$N"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_x$N.json 2> gpurun_out/bench_x$N.err; cat gpurun_out/bench_x$N.json | cut -c1-400; tail -3 gpurun_out/bench_x$N.err
for mode in nccl fused; do
echo "== pagerank $mode x$N scale $CHECK_SCALE (checked)"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/multi_pagerank.py --scale $CHECK_SCALE --edgefactor 16 --iters 20 --mode $mode --check 2> gpurun_out/prc_${mode}_x$N.err | grep '^{' | tee gpurun_out/prc_${mode}_x$N.json
echo "== pagerank $mode x$N scale $SCALE"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/multi_pagerank.py --scale $SCALE --edgefactor 16 --iters 20 --mode $mode 2> gpurun_out/pr_${mode}_x$N.err | grep '^{' | tee gpurun_out/pr_${mode}_x$N.json; tail -2 gpurun_out/pr_${mode}_x$N.err
done
