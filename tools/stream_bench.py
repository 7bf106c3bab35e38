#!/usr/bin/env python
"""The stream schedule on the three operator shapes it serves: RMAT scale 22 (power-law rows), the 8192^2 grid
(4 entries per row, local gathers) and one 64 MB column panel of RMAT scale 26 (short skewed rows).  Tuning aid for
SHB_STREAM_LONG (spmv.cuh: pieces longer than this leave the thread-per-row loop).  One JSON line."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sparseharness_b200 as shb  # noqa: E402


def med(ctx, A, x, out, cfg, reps=12):
    ts = [ctx.spmv(A, shb.PLUS_TIMES, x, None, 1.0, 0.0, out, cfg) for _ in range(reps + 2)][2:]
    return float(np.median(ts)) / 1e6


def main():
    ctx = shb.Context(0)
    res = {k: os.environ.get(k, "default") for k in ("SHB_STREAM_LONG", "SHB_STREAM_SCAN", "SHB_STREAM_TMA")}
    cfg = shb.launch_cfg(variant=shb.VARIANT_STREAM)
    for tag, make in (("config2", lambda: ctx.synth_uniform(1 << 24, 1 << 24, 16)),
                      ("rmat22", lambda: ctx.synth_rmat(22, 16, seed=0x5EED0022, symmetrise=False, value_mode=1)),
                      ("grid8192", lambda: ctx.synth_grid2d(8192, 8192, seed=0x5EED0004))):
        A = make()
        n = A.info.n_rows
        x = ctx.array(np.ones(A.info.n_cols, np.float32))
        out = ctx.empty(n * 4)
        res[tag + "_ms"] = med(ctx, A, x, out, cfg)
        res[tag + "_gbs"] = A.info.algorithmic_bytes_spmv / res[tag + "_ms"] / 1e6
        A.free(), x.free(), out.free()
    print(json.dumps(res))


if __name__ == "__main__":
    main()
