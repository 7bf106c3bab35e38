"""-m gpu: the CUDA SpMV kernels (through the C ABI) against the CPU oracle.

Bit-exact for integer-valued data in all four semirings and for every kernel schedule; bit-exact for
float data when the schedule keeps the gold loop's order (SCALAR); relative 1e-5 (max-norm) for float
data under the reordered schedules (north_star's stated tolerance)."""
import os

import numpy as np
import pytest

import sparseharness_b200 as shb
from conftest import FIXTURES, GOLDEN
from gpu_util import ALL_CFGS, DT, random_ragged, run_spmv, semiring_inputs

pytestmark = pytest.mark.gpu
SEMIRINGS = [shb.PLUS_TIMES, shb.MIN_PLUS, shb.OR_AND, shb.MAX_MIN]
REL_TOL = 1e-5


@pytest.fixture(scope="module")
def ctx():
    c = shb.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("name", list(FIXTURES.values()))
def test_config1_example_matrices_match_reference_gold(ctx, orc, mtx_dir, name):
    """BASELINE config 1: example/matrix.mtx (and the other four fixtures), x=1 y=0 alpha=1 beta=0,
    result == Gold<float>::spmv of the unmodified reference (tests/golden/ref_*.npz), exact `!=` compare
    as Harness::check_result does (inc/harness.h:134)."""
    g = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir[name], "float")
    A = ctx.csr(n, n, rp, col, val)
    for xmode in (0, 1):
        x = orc.xmode_vector(n, xmode)
        for label, cfg in ALL_CFGS:
            got = run_spmv(ctx, A, shb.PLUS_TIMES, x, np.zeros(n, np.float32), 1.0, 0.0, n, **cfg)
            assert np.array_equal(got, g[f"gold_x{xmode}"]), (name, xmode, label)


@pytest.mark.parametrize("name", ["1138_bus", "Trefethen_20", "GD96_b", "arc130", "ca-AstroPh"])
@pytest.mark.parametrize("semiring", SEMIRINGS)
def test_all_semirings_all_schedules_on_fixtures(ctx, orc, mtx_dir, name, semiring):
    elem = "float" if DT[semiring] == np.float32 else "int"
    norm = "scc" if semiring == shb.MAX_MIN else "none"
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir[name], elem, normalise=norm)
    rng = np.random.default_rng(1234 + semiring)
    x, y, alpha, beta = semiring_inputs(orc, semiring, n, n, rng)
    want = orc.spmv(semiring, rp, col, val, x, y, alpha, beta)
    A = ctx.csr(n, n, rp, col, val)
    for label, cfg in ALL_CFGS:
        got = run_spmv(ctx, A, semiring, x, y, alpha, beta, n, **cfg)
        assert np.array_equal(got, want), (name, semiring, label, np.flatnonzero(got != want)[:5])


@pytest.mark.parametrize("semiring", SEMIRINGS)
def test_ragged_rows_long_rows_and_tile_straddlers(ctx, orc, semiring):
    """Unaligned row starts, empty rows, rows longer than one / several stream tiles, n not a multiple of 32."""
    rng = np.random.default_rng(99)
    n, m = 3001, 2777
    rp, col = random_ragged(rng, n, m, mean_len=11, long_rows=[(5, 5000), (6, 4096), (1500, 13001), (2999, 70), (3000, 9000)])
    dt = DT[semiring]
    val = rng.integers(-3, 4, len(col)).astype(dt) if semiring != shb.MIN_PLUS else rng.integers(0, 9, len(col)).astype(dt)
    x, y, alpha, beta = semiring_inputs(orc, semiring, n, m, rng)
    want = orc.spmv(semiring, rp, col, val, x, y, alpha, beta)
    A = ctx.csr(n, m, rp, col, val)
    info = A.info
    assert info.max_row_len == 13001 and info.straddling_rows > 0 and info.stream_tiles == -(-len(col) // 4096)
    for label, cfg in ALL_CFGS:
        got = run_spmv(ctx, A, semiring, x, y, alpha, beta, n, **cfg)
        assert np.array_equal(got, want), (semiring, label, np.flatnonzero(got != want)[:5])
    # pattern operator (no values array): every stored value is the semiring one
    one = {shb.PLUS_TIMES: 1.0, shb.MIN_PLUS: 1.0, shb.OR_AND: 1, shb.MAX_MIN: 2147483647}[semiring]
    wantp = orc.spmv(semiring, rp, col, np.full(len(col), one, dt), x, y, alpha, beta)
    Ap = ctx.csr(n, m, rp, col, None)
    for label, cfg in ALL_CFGS:
        got = run_spmv(ctx, Ap, semiring, x, y, alpha, beta, n, **cfg)
        assert np.array_equal(got, wantp), ("pattern", semiring, label)


@pytest.mark.parametrize("semiring", SEMIRINGS)
def test_stream_merge_row_phase_short_skewed_rows(ctx, orc, semiring):
    """The merge-path row phase of the stream schedule (column panels of power-law operators): mean 2 entries per row,
    a power-law tail, a third of the rows empty, a run of 6000 empty rows (a tile touching more rows than the phase
    holds falls back to the thread-per-row loop), rows spanning threads and tiles.  Both row phases, bit-exact."""
    rng = np.random.default_rng(321)
    n, m = 60001, 5003
    lens = np.minimum((rng.pareto(1.2, n) * 1.5).astype(np.int64), 9000)
    lens[rng.random(n) < 0.33] = 0
    lens[20000:26000] = 0
    lens[[7, 30011, 59999]] = (4096, 10000, 300)
    rp = np.zeros(n + 1, np.int64)
    np.cumsum(lens, out=rp[1:])
    col = rng.integers(0, m, rp[-1]).astype(np.int32)
    dt = DT[semiring]
    val = rng.integers(-3, 4, len(col)).astype(dt) if semiring != shb.MIN_PLUS else rng.integers(0, 9, len(col)).astype(dt)
    x, y, alpha, beta = semiring_inputs(orc, semiring, n, m, rng)
    want = orc.spmv(semiring, rp, col, val, x, y, alpha, beta)
    A = ctx.csr(n, m, rp, col, val)
    for label, flags in (("merge", shb.F_STREAM_MERGE), ("nomerge", shb.F_STREAM_NO_MERGE), ("scan", shb.F_STREAM_SCAN), ("default", 0)):
        got = run_spmv(ctx, A, semiring, x, y, alpha, beta, n, variant=shb.VARIANT_STREAM, flags=flags)
        assert np.array_equal(got, want), (semiring, label, np.flatnonzero(got != want)[:5])


def test_empty_and_degenerate_operators(ctx, orc):
    rng = np.random.default_rng(5)
    # no stored entries at all: out = epi(ZERO, alpha, y, beta)
    n = 70
    rp = np.zeros(n + 1, np.int64)
    A = ctx.csr(n, n, rp, np.zeros(0, np.int32), np.zeros(0, np.float32))
    y = rng.integers(-4, 5, n).astype(np.float32)
    for label, cfg in ALL_CFGS:
        got = run_spmv(ctx, A, shb.PLUS_TIMES, np.ones(n, np.float32), y, 2.0, 3.0, n, **cfg)
        assert np.array_equal(got, 3.0 * y), label
        got = run_spmv(ctx, A, shb.MIN_PLUS, np.ones(n, np.float32), np.abs(y), 0.0, 0.0, n, **cfg)
        assert np.array_equal(got, np.abs(y)), label
    # one row, one column
    A1 = ctx.csr(1, 1, np.array([0, 1]), np.array([0], np.int32), np.array([3.0], np.float32))
    for label, cfg in ALL_CFGS:
        assert run_spmv(ctx, A1, shb.PLUS_TIMES, [2.0], [1.0], 1.0, 1.0, 1, **cfg)[0] == 7.0
    # column indices outside [0, VLength) read ZERO (example/spmv/kernel5.json:3 index guards)
    rp = np.array([0, 3, 6], np.int64)
    col = np.array([0, -1, 5, 1, 7, 2], np.int32)
    val = np.array([1, 10, 100, 2, 20, 4], np.float32)
    x = np.array([1, 1, 1, 1], np.float32)
    A2 = ctx.csr(2, 4, rp, col, val)
    for label, cfg in ALL_CFGS:
        got = run_spmv(ctx, A2, shb.PLUS_TIMES, x, np.zeros(2, np.float32), 1.0, 0.0, 2, **cfg)
        assert list(got) == [1.0, 6.0], label
        got = run_spmv(ctx, A2, shb.MIN_PLUS, x, np.full(2, orc.FLT_MAX, np.float32), 0.0, 0.0, 2, **cfg)
        assert list(got) == [2.0, 3.0], label


def test_float_data_order_and_tolerance(ctx, orc):
    """Non-integer floats: SCALAR keeps the gold loop order -> bit-exact; reversed flag == the ELL
    schedules' order; the reordered schedules are within 1e-5 relative (max-norm)."""
    n = 20000
    rp, col, val = orc.gen_uniform(n, n, 16)
    x = orc.gen_vector(n)
    y = orc.gen_vector(n, seed=77)
    fwd = orc.spmv(orc.PLUS_TIMES, rp, col, val, x, y, 0.75, -1.25, order=0)
    rev = orc.spmv(orc.PLUS_TIMES, rp, col, val, x, y, 0.75, -1.25, order=1)
    A = ctx.csr(n, n, rp, col, val)
    got = run_spmv(ctx, A, shb.PLUS_TIMES, x, y, 0.75, -1.25, n, variant=shb.VARIANT_SCALAR)
    assert np.array_equal(got.view(np.int32), fwd.view(np.int32))
    got = run_spmv(ctx, A, shb.PLUS_TIMES, x, y, 0.75, -1.25, n, variant=shb.VARIANT_SCALAR, flags=shb.F_REVERSED_ORDER)
    assert np.array_equal(got.view(np.int32), rev.view(np.int32))
    scale = np.abs(fwd).max()
    for label, cfg in ALL_CFGS:
        got = run_spmv(ctx, A, shb.PLUS_TIMES, x, y, 0.75, -1.25, n, **cfg)
        assert np.abs(got - fwd).max() <= REL_TOL * scale, label
    # the stream schedule reduces a short row forward from shared memory: also bit-exact here
    got = run_spmv(ctx, A, shb.PLUS_TIMES, x, y, 0.75, -1.25, n, variant=shb.VARIANT_STREAM)
    assert np.array_equal(got.view(np.int32), fwd.view(np.int32))
    # min-plus on float weights is order independent -> exact under every schedule
    w = orc.spmv(orc.MIN_PLUS, rp, col, val, x, y, 0.0, 0.0)
    for label, cfg in ALL_CFGS:
        got = run_spmv(ctx, A, shb.MIN_PLUS, x, y, 0.0, 0.0, n, **cfg)
        assert np.array_equal(got.view(np.int32), w.view(np.int32)), label


def test_y_skipped_only_when_irrelevant(ctx, orc):
    n = 257
    rp, col, val = orc.gen_uniform(n, n, 7, val_mode=1)
    A = ctx.csr(n, n, rp, col, val)
    x = np.ones(n, np.float32)
    want = orc.spmv(orc.PLUS_TIMES, rp, col, val, x, np.zeros(n, np.float32), 1.0, 0.0)
    assert np.array_equal(run_spmv(ctx, A, shb.PLUS_TIMES, x, None, 1.0, 0.0, n), want)
    with pytest.raises(shb.ShbError):
        run_spmv(ctx, A, shb.PLUS_TIMES, x, None, 1.0, 1.0, n)
    with pytest.raises(shb.ShbError):
        run_spmv(ctx, A, shb.MIN_PLUS, x, None, 0.0, 0.0, n)


def test_host_vector_entry_point(ctx, orc):
    n = 5000
    rp, col, val = orc.gen_uniform(n, n, 16, val_mode=1)
    A = ctx.csr(n, n, rp, col, val)
    x = orc.xmode_vector(n, 1)
    y = np.ones(n, np.float32)
    out = np.empty(n, np.float32)
    tot, ker = ctx.spmv_host(A, shb.PLUS_TIMES, x, y, 1.0, 2.0, out)
    assert np.array_equal(out, orc.spmv(orc.PLUS_TIMES, rp, col, val, x, y, 1.0, 2.0)) and tot >= ker > 0


@pytest.mark.parametrize("semiring,cfg", [(shb.PLUS_TIMES, {}), (shb.PLUS_TIMES, dict(variant=shb.VARIANT_SCALAR)),
                                          (shb.MIN_PLUS, dict(variant=shb.VARIANT_VECTOR, lanes=8)),
                                          (shb.PLUS_TIMES, dict(variant=shb.VARIANT_STREAM))])
def test_host_vector_entry_point_large_operator_is_slabbed(ctx, orc, semiring, cfg):
    """>= 2^20 rows: the row-major schedules run in row slabs whose results go back to the host on a second stream
    while the next slab computes; same bits as the device-vector call (ragged last slab, y read at r0 + r)."""
    n = (1 << 20) + 77
    A = ctx.synth_uniform(n, n, 4, val_mode=1)
    x = orc.xmode_vector(n, 1)
    y = (np.arange(n) % 7).astype(np.float32)
    out = np.full(n, -1, np.float32)
    lc = shb.launch_cfg(**cfg)
    tot, ker = ctx.spmv_host(A, semiring, x, y, 1.0, 2.0, out, lc)
    xb, yb, ob = ctx.array(x), ctx.array(y), ctx.empty(n * 4)
    ctx.spmv(A, semiring, xb, yb, 1.0, 2.0, ob, lc)
    assert np.array_equal(out, ob.download(np.float32, n)) and tot >= ker > 0
    rp, col, val = A.download(n - 300, 300)
    want = orc.spmv(semiring, rp, col, val, x, y[n - 300:], 1.0, 2.0)
    assert np.array_equal(out[n - 300:], want)


@pytest.mark.parametrize("n_vec,with_y", [(1, True), (2, False), (5, True), (6, False)])
def test_host_vector_batch_is_pipelined_and_exact(ctx, orc, n_vec, with_y):
    """shb_spmv_host_batch: every vector of the batch gets its own x (and y) and its own result, bit for bit what the
    one-vector call gives, whatever the double-buffer slot it went through."""
    n = 6001
    rp, col, val = orc.gen_uniform(n, n, 9, val_mode=1)
    A = ctx.csr(n, n, rp, col, val)
    rng = np.random.default_rng(n_vec)
    xs = [rng.integers(-4, 5, n).astype(np.float32) for _ in range(n_vec)]
    ys = [rng.integers(-4, 5, n).astype(np.float32) for _ in range(n_vec)] if with_y else None
    outs = [np.full(n, np.nan, np.float32) for _ in range(n_vec)]
    beta = 3.0 if with_y else 0.0
    tot = ctx.spmv_host_batch(A, shb.PLUS_TIMES, xs, ys, 2.0, beta, outs)
    assert tot > 0
    for i in range(n_vec):
        want = orc.spmv(orc.PLUS_TIMES, rp, col, val, xs[i], ys[i] if with_y else np.zeros(n, np.float32), 2.0, beta)
        assert np.array_equal(outs[i], want), i
    if not with_y:
        with pytest.raises(shb.ShbError):
            ctx.spmv_host_batch(A, shb.PLUS_TIMES, xs, None, 2.0, 1.0, outs)  # beta != 0 needs y


def test_argument_validation(ctx):
    A = ctx.csr(4, 4, np.array([0, 1, 2, 3, 4]), np.array([0, 1, 2, 3], np.int32), np.ones(4, np.float32))
    x, out = ctx.array(np.ones(4, np.float32)), ctx.empty(16)
    small = ctx.empty(8)
    with pytest.raises(shb.ShbError):
        ctx.spmv(A, shb.PLUS_TIMES, small, None, 1.0, 0.0, out)
    with pytest.raises(shb.ShbError):
        ctx.spmv(A, shb.PLUS_TIMES, x, None, 1.0, 0.0, x)  # out aliases x
    with pytest.raises(shb.ShbError):
        ctx.spmv(A, 9, x, None, 1.0, 0.0, out)
    with pytest.raises(shb.ShbError):
        ctx.spmv(A, shb.PLUS_TIMES, x, None, 1.0, 0.0, out, shb.launch_cfg(variant=shb.VARIANT_VECTOR, lanes=3))
    with pytest.raises(shb.ShbError):
        ctx.csr(2, 2, np.array([0, 2, 1]), np.array([0], np.int32), None)  # row_ptr must end at nnz


def test_device_generators_equal_oracle_generators(ctx, orc):
    A = ctx.synth_uniform(5000, 4000, 16, row_begin=1000, row_count=3000)
    rp, col, val = A.download()
    orp, ocol, oval = orc.gen_uniform(5000, 4000, 16, row_begin=1000, row_count=3000)
    assert np.array_equal(rp, orp) and np.array_equal(col, ocol) and np.array_equal(val.view(np.int32), oval.view(np.int32))
    for mode in (0, 1):
        assert np.array_equal(ctx.synth_vector(4097, mode=mode).download(np.float32, 4097), orc.gen_vector(4097, mode=mode))
    A = ctx.synth_uniform(300, 300, 5, val_mode=1)
    _, _, v = A.download()
    _, _, ov = orc.gen_uniform(300, 300, 5, val_mode=1)
    assert np.array_equal(v, ov)
    G = ctx.synth_grid2d(37, 23)
    rp, col, val = G.download()
    orp, ocol, oval = orc.gen_grid2d(37, 23)
    assert np.array_equal(rp, orp) and np.array_equal(col, ocol) and np.array_equal(val, oval)
    for sym in (True, False):
        R = ctx.synth_rmat(11, 8, seed=orc.SEED_C3, symmetrise=sym)
        rp, col, _ = R.download()
        orp, ocol = orc.gen_rmat(11, 8, seed=orc.SEED_C3, symmetrise=sym)
        assert np.array_equal(rp, orp) and np.array_equal(col, ocol), sym
    P = ctx.synth_rmat(11, 16, seed=orc.SEED_C5, symmetrise=False, value_mode=1)
    rp, col, val = P.download()
    orp, ocol = orc.gen_rmat(11, 16, seed=orc.SEED_C5, symmetrise=False)
    assert np.array_equal(col, ocol) and np.array_equal(val.view(np.int32), orc.pagerank_values(orp, ocol).view(np.int32))


def test_transpose(ctx, orc):
    rng = np.random.default_rng(3)
    n, m = 700, 450
    rp, col = random_ragged(rng, n, m, mean_len=9, long_rows=[(10, 3000)])
    val = rng.integers(1, 100, len(col)).astype(np.float32)
    A = ctx.csr(n, m, rp, col, val)
    trp, tcol, tval = A.transpose().download()
    rows = np.repeat(np.arange(n), np.diff(rp))
    order = np.lexsort((rows, col))
    assert np.array_equal(tcol, rows[order]) and np.array_equal(tval, val[order])
    assert np.array_equal(trp, np.concatenate([[0], np.cumsum(np.bincount(col, minlength=m))]))


@pytest.mark.parametrize("semiring", SEMIRINGS)
@pytest.mark.parametrize("n_panels", [2, 5, 16])
def test_column_panels_carry_the_row_sums_exactly(ctx, orc, semiring, n_panels):
    """Column panels (shb_csr_build_panels, shb_panels.cu): AUTO runs panel after panel with the row sums carried
    through the epilogue; the result is the one-pass result bit for bit on exact data, in every semiring, with
    y / alpha / beta in play, for one and for several destinations and a row offset (the multi-GPU scatter call)."""
    rng = np.random.default_rng(7 + n_panels)
    n, m = 2500, 3333
    rp, col = random_ragged(rng, n, m, mean_len=9, long_rows=[(3, 6000), (2000, 300)])
    col[rng.random(len(col)) < 0.01] = -1  # out-of-range columns gather ZERO
    dt = DT[semiring]
    val = rng.integers(-3, 4, len(col)).astype(dt) if semiring != shb.MIN_PLUS else rng.integers(0, 9, len(col)).astype(dt)
    x, y, alpha, beta = semiring_inputs(orc, semiring, n, m, rng)
    want = orc.spmv(semiring, rp, col, val, x, y, alpha, beta)
    A = ctx.csr(n, m, rp, col, val)
    assert A.build_panels(n_panels) == n_panels
    got = run_spmv(ctx, A, semiring, x, y, alpha, beta, n)  # AUTO -> panels
    assert np.array_equal(got, want), (semiring, n_panels, np.flatnonzero(got != want)[:5])
    one = run_spmv(ctx, A, semiring, x, y, alpha, beta, n, flags=shb.F_NO_PANELS)
    assert np.array_equal(one, want)
    # scatter form: rows land at an offset in two destinations; y and the convergence test use the offset too
    off = 137
    xb = ctx.array(np.concatenate([x, np.zeros(max(0, off + n - m), dt)]).astype(dt))
    yfull = np.zeros(off + n, dt); yfull[off:] = y
    yb = ctx.array(yfull)
    d0, d1, flag = ctx.empty((off + n) * 4), ctx.empty((off + n) * 4), ctx.empty(4)
    flag.zero()
    ctx.spmv_scatter(A, semiring, xb, yb, alpha, beta, [d0, d1], off, delta=0.0, term=shb.TERM_EQUAL, changed=flag)
    assert np.array_equal(d0.download(dt, n, offset=4 * off), want) and np.array_equal(d1.download(dt, n, offset=4 * off), want)
    A.free()


def test_pagerank_over_column_panels(ctx, orc):
    rp, col = orc.gen_rmat(13, 8, seed=orc.SEED_C3, symmetrise=False)
    val = orc.pagerank_values(rp, col)
    n = len(rp) - 1
    A = ctx.csr(n, n, rp, col, val)
    assert A.build_panels(4) == 4
    res, st = ctx.pagerank(A, delta=1e-7, max_iters=300)
    want, wit = orc.pagerank(rp, col, val, delta=1e-7, max_iters=300, acc_double=True)
    got = res.download(np.float32, n)
    assert st.converged and abs(int(st.iterations) - wit) <= 1
    assert np.abs(got - want).max() <= REL_TOL * np.abs(want).max()
