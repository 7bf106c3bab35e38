#!/usr/bin/env python
"""Single-GPU BFS (config 3, RMAT 24) and SSSP (config 4, 8192^2 grid) a few times each: the command the ncu
launch list / full captures of the graph kernels are taken over (tools/gpu_ncu_graph.sh).  Not a bench."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sparseharness_b200 as shb  # noqa: E402


def main():
    scale = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    side = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    ctx = shb.Context(0)
    A = ctx.synth_rmat(scale, 16, seed=0x5EED0024, symmetrise=True)
    rp = A.download(0, min(4096, 1 << scale))[0]
    root = int(np.argmax(np.diff(rp) > 0))
    for _ in range(reps):
        lv, st = ctx.bfs(A, root, AT=A)
        lv.free()
        print(json.dumps({"bfs_scale": scale, "ms": st.total_ns / 1e6, "levels": st.iterations, "reached": st.reached,
                          "gteps": st.edges_traversed / 2 / st.total_ns}), flush=True)
    A.free()
    G = ctx.synth_grid2d(side, side)
    for _ in range(reps):
        d, st = ctx.sssp(G, 0, AT=G)
        d.free()
        print(json.dumps({"sssp_grid": side, "ms": st.total_ns / 1e6, "rounds": st.iterations,
                          "gteps": st.edges_traversed / st.total_ns}), flush=True)
    G.free()


if __name__ == "__main__":
    main()
