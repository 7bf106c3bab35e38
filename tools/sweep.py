#!/usr/bin/env python
"""GPU-side variant sweep for the SpMV kernels (tuning aid; not part of the product path).

python tools/sweep.py [--rows LOG2] [--reps N]   -> table of ms / effective GB/s per launch configuration
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sparseharness_b200 as shb  # noqa: E402


def time_cfg(ctx, A, sem, x, y, alpha, beta, out, cfg, reps):
    ts = []
    for i in range(reps + 2):
        ns = ctx.spmv(A, sem, x, y, alpha, beta, out, cfg)
        if i >= 2:
            ts.append(ns)
    return float(np.median(ts)) / 1e6, float(np.min(ts)) / 1e6


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=24)
    ap.add_argument("--reps", type=int, default=15)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.json"))
    a = ap.parse_args()
    ctx = shb.Context(0)
    n = 1 << a.rows
    res = []

    def report(tag, A, sem, x, y, alpha, beta, out, cfgs):
        info = A.info
        by = info.algorithmic_bytes_spmv
        for label, kw in cfgs:
            try:
                med, best = time_cfg(ctx, A, sem, x, y, alpha, beta, out, shb.launch_cfg(**kw), a.reps)
                row = {"workload": tag, "cfg": label, "ms_median": med, "ms_best": best, "gbs_median": by / med / 1e6,
                       "gbs_best": by / best / 1e6, "bytes": by}
            except Exception as e:
                row = {"workload": tag, "cfg": label, "error": repr(e)[:200]}
            res.append(row)
            print(json.dumps(row), flush=True)

    V, S, T = shb.VARIANT_VECTOR, shb.VARIANT_SCALAR, shb.VARIANT_STREAM
    NH = shb.F_NO_CACHE_HINTS
    cfgs = [("auto", {}),
            ("vec L4 R4", dict(variant=V, lanes=4, rows_per_group=4)), ("vec L4 R2", dict(variant=V, lanes=4, rows_per_group=2)),
            ("vec L4 R1", dict(variant=V, lanes=4, rows_per_group=1)), ("vec L4 R4 nohint", dict(variant=V, lanes=4, rows_per_group=4, flags=NH)),
            ("vec L4 R4 b128", dict(variant=V, lanes=4, rows_per_group=4, block_threads=128)),
            ("vec L4 R4 b64", dict(variant=V, lanes=4, rows_per_group=4, block_threads=64)),
            ("vec L2 R2", dict(variant=V, lanes=2)), ("vec L8 R2", dict(variant=V, lanes=8, rows_per_group=2)),
            ("vec L8 R4", dict(variant=V, lanes=8, rows_per_group=4)), ("vec L16 R2", dict(variant=V, lanes=16)),
            ("vec L1 R2", dict(variant=V, lanes=1, rows_per_group=2)),
            ("stream", dict(variant=T)), ("stream nohint", dict(variant=T, flags=NH)), ("scalar", dict(variant=S))]

    # config 2: uniform random 16 nnz/row, float plus-times
    A = ctx.synth_uniform(n, n, 16)
    x = ctx.synth_vector(n)
    out = ctx.empty(n * 4)
    report(f"uniform 2^{a.rows} x16 f32 plus-times", A, shb.PLUS_TIMES, x, None, 1.0, 0.0, out, cfgs)
    report(f"uniform 2^{a.rows} x16 f32 min-plus", A, shb.MIN_PLUS, x, x, 0.0, 0.0, out, cfgs[:2] + cfgs[12:13])
    A.free()
    # same structure, x gathered from a 1 MB window only (columns < 2^18): isolates the gather cost
    A = ctx.synth_uniform(n, 1 << 18, 16)
    report(f"uniform 2^{a.rows} x16, cols < 2^18 (L2/L1-friendly gather)", A, shb.PLUS_TIMES, x, None, 1.0, 0.0, out,
           cfgs[:2] + cfgs[12:13])
    A.free()
    # config 4 family: 2-D grid sweep (local gathers): the HBM-bound case
    side = 1 << (a.rows // 2)
    G = ctx.synth_grid2d(side, (1 << a.rows) // side)
    report(f"grid {side}x{(1 << a.rows) // side} f32 min-plus sweep", G, shb.MIN_PLUS, x, x, 0.0, 0.0, out,
           [("auto", {}), ("vec L1 R2", dict(variant=V, lanes=1, rows_per_group=2)), ("vec L2 R2", dict(variant=V, lanes=2)),
            ("vec L4 R4", dict(variant=V, lanes=4, rows_per_group=4)), ("stream", dict(variant=T)), ("scalar", dict(variant=S))])
    G.free()
    # config 3/5 family: RMAT (power-law rows)
    sc = max(a.rows - 2, 10)
    R = ctx.synth_rmat(sc, 16, seed=0x5EED0026, symmetrise=False, value_mode=1)
    info = R.info
    print(json.dumps({"rmat_scale": sc, "nnz": info.nnz, "max_row": info.max_row_len, "auto": [info.auto_variant, info.auto_lanes]}))
    xr = ctx.synth_vector(1 << sc)
    outr = ctx.empty((1 << sc) * 4)
    report(f"rmat s{sc} ef16 directed f32 plus-times", R, shb.PLUS_TIMES, xr, None, 1.0, 0.0, outr,
           [("auto", {}), ("stream", dict(variant=T)), ("vec L4 R4", dict(variant=V, lanes=4, rows_per_group=4)),
            ("vec L8 R2", dict(variant=V, lanes=8)), ("vec L16 R2", dict(variant=V, lanes=16)), ("vec L32 R1", dict(variant=V, lanes=32)),
            ("scalar", dict(variant=S))])
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(res, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
