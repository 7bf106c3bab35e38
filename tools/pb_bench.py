#!/usr/bin/env python
"""Times the BLOCKED schedule (pb.cuh) on BASELINE config 2 against the row-major schedules (tuning aid).

python tools/pb_bench.py [--rows LOG2] [--rs 4096,2048] [--reps N] [--rmat SCALE]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sparseharness_b200 as shb  # noqa: E402


def med(ctx, A, sem, x, y, alpha, beta, out, cfg, reps):
    ts = [ctx.spmv(A, sem, x, y, alpha, beta, out, cfg) for _ in range(reps + 3)][3:]
    return float(np.median(ts)) / 1e6


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=24)
    ap.add_argument("--rs", default="4096")
    ap.add_argument("--reps", type=int, default=15)
    ap.add_argument("--rmat", type=int, default=0)
    ap.add_argument("--only-blocked", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    a = ap.parse_args()
    ctx = shb.Context(0)
    n = 1 << a.rows
    A = ctx.synth_uniform(n, n, 16)
    x = ctx.synth_vector(n)
    out, ref = ctx.empty(n * 4), ctx.empty(n * 4)
    by = A.info.algorithmic_bytes_spmv
    t = med(ctx, A, shb.PLUS_TIMES, x, None, 1.0, 0.0, ref, shb.launch_cfg(variant=shb.VARIANT_VECTOR, lanes=4),
            2 if a.only_blocked else a.reps)
    print(json.dumps({"cfg": "vector L4", "ms": t, "gbs": by / t / 1e6}), flush=True)
    want = ref.download(np.float32, n)
    for rs in [int(v) for v in a.rs.split(",")]:
        geom = A.build_blocked(rs, make_default=False)
        for sem, name in ((shb.PLUS_TIMES, "plus-times"), (shb.MIN_PLUS, "min-plus")):
            t = med(ctx, A, sem, x, x if sem == shb.MIN_PLUS else None, 1.0 if sem == shb.PLUS_TIMES else 0.0, 0.0, out,
                    shb.launch_cfg(variant=shb.VARIANT_BLOCKED), a.reps)
            row = {"cfg": f"blocked rs={rs} {name}", "ms": t, "gbs": by / t / 1e6, "geom": geom}
            if sem == shb.PLUS_TIMES and not a.no_check:
                got = out.download(np.float32, n)
                row["max_rel_err_vs_vector"] = float(np.abs(got - want).max() / np.abs(want).max())
            print(json.dumps(row), flush=True)
    A.free()
    if a.rmat:
        R = ctx.synth_rmat(a.rmat, 16, seed=0x5EED0026, symmetrise=False, value_mode=1)
        m = 1 << a.rmat
        xr, o1, o2 = ctx.synth_vector(m), ctx.empty(m * 4), ctx.empty(m * 4)
        byr = R.info.algorithmic_bytes_spmv
        t = med(ctx, R, shb.PLUS_TIMES, xr, None, 1.0, 0.0, o1, shb.launch_cfg(), a.reps)
        print(json.dumps({"cfg": f"rmat s{a.rmat} auto", "ms": t, "gbs": byr / t / 1e6}), flush=True)
        geom = R.build_blocked(0, make_default=False)
        t = med(ctx, R, shb.PLUS_TIMES, xr, None, 1.0, 0.0, o2, shb.launch_cfg(variant=shb.VARIANT_BLOCKED), a.reps)
        w, g = o1.download(np.float32, m), o2.download(np.float32, m)
        print(json.dumps({"cfg": f"rmat s{a.rmat} blocked", "ms": t, "gbs": byr / t / 1e6, "geom": geom,
                          "max_rel_err": float(np.abs(g - w).max() / np.abs(w).max())}), flush=True)


if __name__ == "__main__":
    main()
