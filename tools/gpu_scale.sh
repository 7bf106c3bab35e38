#!/bin/bash
# bench.py under torchrun at N GPUs (what the driver's scaling run does), output kept under gpurun_out/
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 5 > gpurun_out/bench_x$N.json 2> gpurun_out/bench_x$N.err
echo "rc=$?"; cat gpurun_out/bench_x$N.json; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/bench_x$N.err | tail -5
