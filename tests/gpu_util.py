"""Helpers shared by the -m gpu parity tests: everything device-side goes through the C ABI
(sparseharness_b200 -> lib/libshb.so); the oracle is only the checker."""
import numpy as np

import sparseharness_b200 as shb

DT = {shb.PLUS_TIMES: np.float32, shb.MIN_PLUS: np.float32, shb.OR_AND: np.int32, shb.MAX_MIN: np.int32}

ALL_CFGS = [("scalar", dict(variant=shb.VARIANT_SCALAR)), ("stream", dict(variant=shb.VARIANT_STREAM)),
            ("auto", dict(variant=shb.VARIANT_AUTO))] + \
           [(f"vector{l}", dict(variant=shb.VARIANT_VECTOR, lanes=l)) for l in (1, 2, 4, 8, 16, 32)] + \
           [("vector4r1", dict(variant=shb.VARIANT_VECTOR, lanes=4, rows_per_group=1)),
            ("vector8r4", dict(variant=shb.VARIANT_VECTOR, lanes=8, rows_per_group=4)),
            ("vector4nohint", dict(variant=shb.VARIANT_VECTOR, lanes=4, flags=shb.F_NO_CACHE_HINTS)),
            ("stream_nohint", dict(variant=shb.VARIANT_STREAM, flags=shb.F_NO_CACHE_HINTS)),
            ("stream_merge", dict(variant=shb.VARIANT_STREAM, flags=shb.F_STREAM_MERGE)),        # merge-path row phase
            ("stream_nomerge", dict(variant=shb.VARIANT_STREAM, flags=shb.F_STREAM_NO_MERGE)),   # thread-per-row row phase
            ("stream_scan", dict(variant=shb.VARIANT_STREAM, flags=shb.F_STREAM_SCAN)),          # segmented-scan row phase
            ("stream_tma", dict(variant=shb.VARIANT_STREAM, flags=shb.F_STREAM_TMA)),            # tile streams by TMA bulk copies
            ("blocked", dict(variant=shb.VARIANT_BLOCKED)),          # one row panel on the small fixtures
            ("blocked_rs5", dict(variant=shb.VARIANT_BLOCKED, pb_rows=5))]  # many row panels, tiny pieces


def run_spmv(ctx, A, semiring, x, y, alpha, beta, n_rows, **cfg):
    dt = DT[semiring]
    cfg = dict(cfg)
    pb_rows = cfg.pop("pb_rows", 0)
    if cfg.get("variant") == shb.VARIANT_BLOCKED and A.info.nnz > 0 and getattr(A, "_pb_rows", None) != pb_rows:
        A.build_blocked(pb_rows, make_default=False)  # the operator's blocked copy (pb.cuh), rebuilt per geometry
        A._pb_rows = pb_rows
    xb = ctx.array(np.asarray(x, dt))
    yb = ctx.array(np.asarray(y, dt)) if y is not None else None
    out = ctx.empty(max(n_rows, 1) * 4)
    out.fill_u32(0xDEADBEEF)
    ctx.spmv(A, semiring, xb, yb, alpha, beta, out, shb.launch_cfg(**cfg))
    return out.download(dt, n_rows)


def semiring_inputs(orc, semiring, n_rows, n_cols, rng):
    """(x, y, alpha, beta) with small-integer content so that every semiring result is exact."""
    if semiring == shb.PLUS_TIMES:
        return (rng.integers(-5, 6, n_cols).astype(np.float32), rng.integers(-4, 5, n_rows).astype(np.float32), 2.0, 3.0)
    if semiring == shb.MIN_PLUS:
        x = rng.integers(0, 60, n_cols).astype(np.float32)
        x[rng.random(n_cols) < 0.3] = orc.FLT_MAX
        y = rng.integers(0, 60, n_rows).astype(np.float32)
        y[rng.random(n_rows) < 0.3] = orc.FLT_MAX
        return x, y, 1.0, 2.0
    if semiring == shb.OR_AND:
        return ((rng.random(n_cols) < 0.2).astype(np.int32) * rng.integers(1, 9, n_cols).astype(np.int32),
                (rng.random(n_rows) < 0.3).astype(np.int32) * 5, 1, 1)
    x = rng.permutation(n_cols).astype(np.int32)
    y = rng.integers(-50, n_cols, n_rows).astype(np.int32)
    return x, y, int(n_cols // 2), 7


def random_ragged(rng, n_rows, n_cols, mean_len=12, long_rows=(), empty_frac=0.1):
    lens = rng.geometric(1.0 / mean_len, n_rows).astype(np.int64)
    lens[rng.random(n_rows) < empty_frac] = 0
    for r, l in long_rows:
        lens[r] = l
    rp = np.zeros(n_rows + 1, np.int64)
    np.cumsum(lens, out=rp[1:])
    col = rng.integers(0, n_cols, rp[-1]).astype(np.int32)
    return rp, col
