#!/bin/bash
# BFS check: parity tests, per-level trace of the partitioned code on one GPU, the single-GPU driver
mkdir -p gpurun_out
echo "== pytest bfs"; timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_iterate.py tests/test_gpu_fullsize.py -m gpu -x -q --timeout 600 -k "bfs or config3" 2>&1 | tail -4
echo "== bfs x1 s24 trace"; timeout 600 python tools/multi_bfs.py --scale 24 --trace 2>gpurun_out/t1.err | tee gpurun_out/bfs_trace_x1.json | cut -c700-1600; tail -2 gpurun_out/t1.err
echo "== single-GPU shb_bfs s24"; timeout 600 python - <<'PY'
import numpy as np, sparseharness_b200 as shb
ctx = shb.Context(0)
A = ctx.synth_rmat(24, 16, seed=0x5EED0024, symmetrise=True)
rp = A.download(0, 4096)[0]; root = int(np.argmax(np.diff(rp) > 0))
for _ in range(3):
    lv, st = ctx.bfs(A, root, AT=A); lv.free()
    print("shb_bfs s24: %.3f ms, %d levels, %.1f GTEPS" % (st.total_ns / 1e6, st.iterations, st.edges_traversed / 2 / st.total_ns))
PY
