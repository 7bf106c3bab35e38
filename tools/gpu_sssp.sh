#!/bin/bash
# SSSP tuning call: parity tests, then config 4 (8192^2 grid) under a few bucket widths.
mkdir -p gpurun_out
echo "== pytest sssp"; timeout 900 python -m pytest tests/test_gpu_iterate.py tests/test_gpu_fullsize.py -m gpu -x -q --timeout 600 -k "sssp" 2>&1 | tail -15
cat > /tmp/sssp_run.py <<'PY'
import os, sys, time
sys.path.insert(0, os.getcwd())
import sparseharness_b200 as shb
ctx = shb.Context(0)
G = ctx.synth_grid2d(8192, 8192)
for rep in range(2):
    dist, st = ctx.sssp(G, 0, AT=G)
    print("grid8192", os.environ.get("SHB_SSSP_C"), os.environ.get("SHB_SSSP_CTAS_PER_SM"), st.iterations, st.converged, st.total_ns / 1e6, "ms", flush=True)
PY
export SHB_SSSP_VERBOSE=1
for C in 32 64 128 256; do SHB_SSSP_C=$C timeout 300 python /tmp/sssp_run.py 2>&1 | tail -4; done
