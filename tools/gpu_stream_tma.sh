#!/bin/bash
# TMA-staged phase 1 of the stream schedule: parity, then A/B on config 2 / RMAT-22 / grid
timeout 300 python -m pytest tests/test_gpu_spmv.py -m gpu -x -q --timeout 120 2>&1 | tail -3
for v in 0 1; do SHB_STREAM_TMA=$v timeout 200 python tools/stream_bench.py 2>&1 | tail -1; done
