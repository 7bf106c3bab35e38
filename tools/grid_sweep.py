#!/usr/bin/env python
"""Dense min-plus sweep over the config-4 grid operator under several schedules (tuning aid; the HBM-bound case)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sparseharness_b200 as shb  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
only = sys.argv[2] if len(sys.argv) > 2 else None
ctx = shb.Context(0)
G = ctx.synth_grid2d(side, side)
info = G.info
n = info.n_rows
x = ctx.synth_vector(n)
out = ctx.empty(n * 4)
V, T, S = shb.VARIANT_VECTOR, shb.VARIANT_STREAM, shb.VARIANT_SCALAR
cfgs = [("auto", {}), ("stream", dict(variant=T)), ("vec L1 R2", dict(variant=V, lanes=1, rows_per_group=2)),
        ("vec L1 R4", dict(variant=V, lanes=1, rows_per_group=4)), ("vec L2 R2", dict(variant=V, lanes=2, rows_per_group=2)),
        ("vec L2 R4", dict(variant=V, lanes=2, rows_per_group=4)), ("vec L4 R2", dict(variant=V, lanes=4, rows_per_group=2)),
        ("vec L4 R4", dict(variant=V, lanes=4, rows_per_group=4)), ("scalar", dict(variant=S))]
for label, kw in cfgs:
    if only and label != only:
        continue
    for sem, name in ((shb.MIN_PLUS, "min-plus"), (shb.PLUS_TIMES, "plus-times")):
        ts = [ctx.spmv(G, sem, x, x, 0.0 if sem == shb.MIN_PLUS else 1.0, 0.0, out, shb.launch_cfg(**kw)) for _ in range(12)][2:]
        ms = float(np.median(ts)) / 1e6
        print(json.dumps({"grid": side, "cfg": label, "semiring": name, "ms": ms,
                          "gbs": info.algorithmic_bytes_spmv / ms / 1e6, "auto": [info.auto_variant, info.auto_lanes]}), flush=True)
