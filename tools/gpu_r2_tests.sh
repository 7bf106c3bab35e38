#!/bin/bash
# Round 2, call 1: the whole -m gpu suite (new pinned + full-size oracle tests), then a short bench with the parity gates.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/nvsmi.txt 2>&1
nproc > gpurun_out/nproc.txt; free -g >> gpurun_out/nproc.txt
echo "== pytest gpu"; ( time timeout 1700 python -m pytest tests -m gpu -x -q --timeout 900 --durations=15 ) > gpurun_out/r2_pytest.log 2>&1; tail -40 gpurun_out/r2_pytest.log
echo "== bench"; ( time timeout 900 python bench.py --steps 50 --warmup 5 ) > gpurun_out/r2_bench1.json 2> gpurun_out/r2_bench1.err; cat gpurun_out/r2_bench1.json | cut -c1-3000; tail -8 gpurun_out/r2_bench1.err
