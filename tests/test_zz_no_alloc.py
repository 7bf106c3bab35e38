"""shb_spmv_stream_static (include/shb.h): the stream-schedule SpMV whose scratch is the library's file-scope __device__
array, for callers that cannot allocate.  No counterpart in the reference.

Checked through the C ABI against the existing entry point (shb_spmv with the stream schedule): the two run the same
kernels, so the results are compared bit for bit -- which is inside every tolerance the other tests use (exact for the
integer semirings and integer-valued floats, 1e-5 relative for general floats) -- on two calls in a row that share the
static array (the second on a different operator and semiring, so the array still holds the first call's carries), then
again after both, at the size limit, and just above it (rejected with a clear error).

The file's name keeps it last in pytest's collection order."""
import numpy as np
import pytest

import sparseharness_b200 as shb
from gpu_util import DT, random_ragged, semiring_inputs

pytestmark = pytest.mark.gpu

STREAM = dict(variant=shb.VARIANT_STREAM)


@pytest.fixture(scope="module")
def ctx():
    c = shb.Context(0)
    yield c
    c.close()


def _both(ctx, A, semiring, x, y, alpha, beta, n):
    """(existing entry point with the stream schedule, the no-allocation entry point) on the same device buffers"""
    dt = DT[semiring]
    xb = ctx.array(np.asarray(x, dt))
    yb = ctx.array(np.asarray(y, dt)) if y is not None else None
    ref, got = ctx.empty(max(n, 1) * 4), ctx.empty(max(n, 1) * 4)
    ref.fill_u32(0xDEADBEEF)
    got.fill_u32(0xDEADBEEF)
    ctx.spmv(A, semiring, xb, yb, alpha, beta, ref, shb.launch_cfg(**STREAM))
    ctx.spmv_stream_static(A, semiring, xb, yb, alpha, beta, got)
    r, g = ref.download(dt, n), got.download(dt, n)
    for b in (xb, yb, ref, got):
        if b is not None:
            b.free()
    return r, g


def _ragged_operator(ctx, rng, semiring, n, m, long_rows):
    rp, col = random_ragged(rng, n, m, mean_len=11, long_rows=long_rows)
    dt = DT[semiring]
    val = rng.integers(-3, 4, len(col)).astype(dt) if semiring != shb.MIN_PLUS else rng.integers(0, 9, len(col)).astype(dt)
    return ctx.csr(n, m, rp, col, val), rp, col, val


def test_two_calls_in_a_row_match_the_existing_entry_point(ctx, orc):
    rng = np.random.default_rng(2024)
    # call 1: plus-times, rows that straddle one and several tiles (the carries live in the static array)
    n1, m1 = 3001, 2777
    A1, rp1, col1, val1 = _ragged_operator(ctx, rng, shb.PLUS_TIMES, n1, m1, [(5, 5000), (6, 4096), (1500, 13001), (3000, 9000)])
    assert A1.info.straddling_rows > 0
    x1, y1, a1, b1 = semiring_inputs(orc, shb.PLUS_TIMES, n1, m1, rng)
    ref1, got1 = _both(ctx, A1, shb.PLUS_TIMES, x1, y1, a1, b1, n1)
    assert np.array_equal(ref1.view(np.uint32), got1.view(np.uint32))
    assert np.array_equal(got1, orc.spmv(shb.PLUS_TIMES, rp1, col1, val1, x1, y1, a1, b1))  # and both equal the oracle
    # call 2, right after, another operator and semiring: fewer tiles, the array still holds call 1's carries
    n2, m2 = 1203, 901
    A2, rp2, col2, val2 = _ragged_operator(ctx, rng, shb.MIN_PLUS, n2, m2, [(0, 4500), (700, 8200)])
    x2, y2, a2, b2 = semiring_inputs(orc, shb.MIN_PLUS, n2, m2, rng)
    ref2, got2 = _both(ctx, A2, shb.MIN_PLUS, x2, y2, a2, b2, n2)
    assert np.array_equal(ref2.view(np.uint32), got2.view(np.uint32))
    assert np.array_equal(got2, orc.spmv(shb.MIN_PLUS, rp2, col2, val2, x2, y2, a2, b2))
    # call 3: the first operator again (general float values: the 1e-5 case of the other tests; same kernels, so exact)
    xf = rng.uniform(-1, 1, m1).astype(np.float32)
    ref3, got3 = _both(ctx, A1, shb.PLUS_TIMES, xf, None, 1.0, 0.0, n1)
    assert np.array_equal(ref3.view(np.uint32), got3.view(np.uint32))
    # integer semirings and a pattern operator (no values array)
    for sem in (shb.OR_AND, shb.MAX_MIN):
        A, rp, col, val = _ragged_operator(ctx, rng, sem, n2, m2, [(3, 6000)])
        x, y, al, be = semiring_inputs(orc, sem, n2, m2, rng)
        r, g = _both(ctx, A, sem, x, y, al, be, n2)
        assert np.array_equal(r, g) and np.array_equal(g, orc.spmv(sem, rp, col, val, x, y, al, be)), sem
        A.free()
    Ap = ctx.csr(n2, m2, rp2, col2, None)
    r, g = _both(ctx, Ap, shb.PLUS_TIMES, x1[:m2], None, 1.0, 0.0, n2)
    assert np.array_equal(r.view(np.uint32), g.view(np.uint32))
    # no stored entries: out = epi(ZERO, alpha, y, beta), no tiles, no scratch
    A0 = ctx.csr(70, 70, np.zeros(71, np.int64), np.zeros(0, np.int32), np.zeros(0, np.float32))
    y0 = rng.integers(-4, 5, 70).astype(np.float32)
    r, g = _both(ctx, A0, shb.PLUS_TIMES, np.ones(70, np.float32), y0, 2.0, 3.0, 70)
    assert np.array_equal(r, g) and np.array_equal(g, 3.0 * y0)
    for A in (A1, A2, Ap, A0):
        A.free()


def test_size_limit_is_stated_and_enforced(ctx):
    import ctypes as C

    from sparseharness_b200._capi import lib
    lim = C.c_int64(0)
    assert lib.shb_spmv_stream_static_max_nnz(C.byref(lim)) == 0 and lim.value == 1 << 28
    # exactly at the limit: BASELINE config 2's operator (2^24 rows x 16 = 2^28 stored entries = 65536 tiles)
    n = 1 << 24
    A = ctx.synth_uniform(n, n, 16, val_mode=1)
    x = ctx.synth_vector(n, mode=1)
    ref, got = ctx.empty(n * 4), ctx.empty(n * 4)
    ctx.spmv(A, shb.PLUS_TIMES, x, None, 1.0, 0.0, ref, shb.launch_cfg(**STREAM))
    ctx.spmv_stream_static(A, shb.PLUS_TIMES, x, None, 1.0, 0.0, got)
    assert np.array_equal(ref.download(np.uint32, n), got.download(np.uint32, n))
    A.free()
    # one tile more: rejected, with the limit in the message, and nothing launched
    nb = n + 256
    B = ctx.synth_uniform(nb, n, 16, val_mode=1)
    assert B.info.nnz > lim.value
    out = ctx.empty(nb * 4)
    with pytest.raises(shb.ShbError) as e:
        ctx.spmv_stream_static(B, shb.PLUS_TIMES, x, None, 1.0, 0.0, out)
    assert e.value.status == 6 and "65536 tiles" in str(e.value)  # SHB_ERR_TOO_LARGE
    ctx.spmv(B, shb.PLUS_TIMES, x, None, 1.0, 0.0, out, shb.launch_cfg(**STREAM))  # the regular entry point still serves it
    for b in (B, x, ref, got, out):
        b.free()
