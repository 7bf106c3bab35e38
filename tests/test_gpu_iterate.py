"""-m gpu: device-resident fixed-point loops and graph drivers against the CPU oracle (bit-exact for
BFS levels, min-plus distances, max-min labels; 1e-5 relative for PageRank)."""
import numpy as np
import pytest

import sparseharness_b200 as shb
from gpu_util import ALL_CFGS, random_ragged

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = shb.Context(0)
    yield c
    c.close()


def _iterate(ctx, A, semiring, x0, y0, alpha, beta, dt, **kw):
    n = len(x0)
    xb, yb, wb = ctx.array(x0.astype(dt)), ctx.array(y0.astype(dt)), ctx.empty(n * 4)
    wb.zero()
    res, iters, total_ns, per, ok = ctx.iterate(A, semiring, xb, yb, alpha, beta, wb, **kw)
    return res.download(dt, n), iters, per, ok


def test_sssp_example_faithful_iteration(ctx, orc, mtx_dir):
    """scripts/example/sssp.sh: sssp_harness on matrix3, x = y = (0, FLT_MAX...), alpha = beta = 0,
    loop of app/sssp.cpp:112-153 with y re-bound to the input; same result AND same launch count."""
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir["Trefethen_20"], "float")
    x0 = np.full(n, orc.FLT_MAX, np.float32); x0[0] = 0
    want, want_iters = orc.iterate(orc.MIN_PLUS, rp, col, val, x0, x0.copy(), 0.0, 0.0, delta=1e-4)
    A = ctx.csr(n, n, rp, col, val)
    for check_every in (1, 3, 16):
        for label, cfg in ALL_CFGS[:6]:
            got, iters, per, ok = _iterate(ctx, A, shb.MIN_PLUS, x0, x0.copy(), 0.0, 0.0, np.float32, delta=1e-4,
                                           check_every=check_every, per_iter=True, cfg=shb.launch_cfg(**cfg))
            assert ok and iters == want_iters and len(per) == iters and (per > 0).all(), (label, check_every)
            assert np.array_equal(got, want), (label, check_every)


@pytest.mark.parametrize("name", ["1138_bus", "GD96_b", "arc130", "ca-AstroPh"])
def test_faithful_loops_all_apps(ctx, orc, mtx_dir, name):
    cap = 40
    # bfs (app/bfs.cpp:209-216: x = y = e_0, alpha 1, beta 0; stop on equality; may never stop -> cap, Q8)
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir[name], "int")
    x0 = np.zeros(n, np.int32); x0[0] = 1
    want, wi = orc.iterate(orc.OR_AND, rp, col, val, x0, x0.copy(), 1, 0, max_iters=cap)
    A = ctx.csr(n, n, rp, col, val)
    got, gi, _, ok = _iterate(ctx, A, shb.OR_AND, x0, x0.copy(), 1, 0, np.int32, max_iters=cap, check_every=7)
    assert gi == wi and np.array_equal(got, want) and (ok or wi == cap)
    # scc (app/scc.cpp:197-222: x[i] = i, y = INT_MIN, alpha INT_MAX, beta INT_MIN, scc_normalise)
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir[name], "int", normalise="scc")
    x0 = np.arange(n, dtype=np.int32); y0 = np.full(n, orc.INT_MIN, np.int32)
    want, wi = orc.iterate(orc.MAX_MIN, rp, col, val, x0, y0, orc.INT_MAX, orc.INT_MIN, max_iters=cap)
    A = ctx.csr(n, n, rp, col, val)
    got, gi, _, _ = _iterate(ctx, A, shb.MAX_MIN, x0, y0, int(orc.INT_MAX), int(orc.INT_MIN), np.int32, max_iters=cap)
    assert gi == wi and np.array_equal(got, want)
    # pr (app/pr.cpp:182-201: x = 1/N, y = 1, alpha 1, beta (1-d)/N, pagerank_normalise then Q2 truncation, y re-bound)
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir[name], "float", normalise="pr")
    x0 = np.full(n, np.float32(1.0) / np.float32(n), np.float32); y0 = np.ones(n, np.float32)
    beta = (np.float32(1.0) - np.float32(0.85)) / np.float32(n)
    want, wi = orc.iterate(orc.PLUS_TIMES, rp, col, val, x0, y0, 1.0, beta, delta=1e-4, max_iters=cap)
    A = ctx.csr(n, n, rp, col, val)
    got, gi, _, _ = _iterate(ctx, A, shb.PLUS_TIMES, x0, y0, 1.0, float(beta), np.float32, delta=1e-4, max_iters=cap,
                             cfg=shb.launch_cfg(variant=shb.VARIANT_SCALAR))
    # (on several fixtures the reference recurrence diverges to inf/NaN -- 'non-working' per TODO.md; the
    #  device must diverge identically: same launch count, same values, NaN where the CPU has NaN)
    assert gi == wi and np.array_equal(got, want, equal_nan=True)
    fin = np.isfinite(want)
    assert np.array_equal(got[fin].view(np.int32), want[fin].view(np.int32))
    # sssp
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir[name], "float")
    x0 = np.full(n, orc.FLT_MAX, np.float32); x0[0] = 0
    want, wi = orc.iterate(orc.MIN_PLUS, rp, col, val, x0, x0.copy(), 0.0, 0.0, delta=1e-4, max_iters=400)
    A = ctx.csr(n, n, rp, col, val)
    got, gi, _, _ = _iterate(ctx, A, shb.MIN_PLUS, x0, x0.copy(), 0.0, 0.0, np.float32, delta=1e-4, max_iters=400)
    assert gi == wi and np.array_equal(got, want)
    # frontier-aware SSSP reaches the same fixed point as the dense sweeps and as Dijkstra, bit for bit
    dist, st = ctx.sssp(A, 0)
    d = dist.download(np.float32, n)
    assert st.converged and np.array_equal(d, orc.sssp(rp, col, val, 0))
    if wi < 400:
        assert np.array_equal(d, want)


def test_iteration_cap_reports_not_converged(ctx, orc):
    # 2-cycle: or-and walk parity never settles (Q8 / TODO.md:7-9) -> cap, status NOT_CONVERGED, no hang
    rp = np.array([0, 1, 2], np.int64); col = np.array([1, 0], np.int32); val = np.ones(2, np.int32)
    A = ctx.csr(2, 2, rp, col, val)
    x0 = np.array([1, 0], np.int32)
    got, iters, _, ok = _iterate(ctx, A, shb.OR_AND, x0, x0.copy(), 1, 0, np.int32, max_iters=11, check_every=4)
    assert not ok and iters == 11
    want, wi = orc.iterate(orc.OR_AND, rp, col, val, x0, x0.copy(), 1, 0, max_iters=11)
    assert wi == 11 and np.array_equal(got, want)


@pytest.mark.parametrize("with_transpose", [False, True])
@pytest.mark.parametrize("scale,ef", [(8, 4), (13, 16)])
def test_bfs_levels_rmat(ctx, orc, scale, ef, with_transpose):
    A = ctx.synth_rmat(scale, ef, seed=orc.SEED_C3, symmetrise=True)
    rp, col, _ = A.download()
    deg = np.diff(rp)
    root = int(np.argmax(deg > 0))
    want, nlev, edges = orc.bfs_levels(rp, col, None, root)
    levels, st = ctx.bfs(A, root, AT=A if with_transpose else None)  # symmetric: A is its own transpose
    got = levels.download(np.int32, len(rp) - 1)
    assert np.array_equal(got, want)
    assert st.converged and st.iterations == nlev and st.reached == int((want >= 0).sum()) and st.edges_traversed == edges


def test_bfs_directed_with_zero_valued_edges(ctx, orc):
    rng = np.random.default_rng(8)
    n = 1500
    rp, col = random_ragged(rng, n, n, mean_len=3, long_rows=[(7, 900)], empty_frac=0.2)
    val = (rng.random(len(col)) < 0.8).astype(np.int32) * rng.integers(1, 5, len(col)).astype(np.int32)
    A = ctx.csr(n, n, rp, col, val)
    want, nlev, edges = orc.bfs_levels(rp, col, val, 3)
    for AT in (None, A.transpose()):
        levels, st = ctx.bfs(A, 3, AT=AT)
        assert np.array_equal(levels.download(np.int32, n), want) and st.iterations == nlev


def test_bfs_more_hub_vertices_in_a_level_than_the_hub_list_holds(ctx, orc):
    """Push levels park vertices with >= 1024 out-edges in a list of 1024 for the whole grid; a level with more hubs than
    that falls back to the warp expansion for the rest.  Symmetric graph: 3000 vertices with ~1400 random neighbours each,
    so level 1 is ~1400 hubs.  Also: no edges at all, and a root without edges."""
    rng = np.random.default_rng(41)
    n, k = 3000, 800
    src = np.repeat(np.arange(n, dtype=np.int64), k)
    dst = rng.integers(0, n, n * k, dtype=np.int64)
    key = np.unique(np.concatenate([src * n + dst, dst * n + src]))
    key = key[(key // n) != (key % n)]
    rows, col = key // n, (key % n).astype(np.int32)
    rp = np.zeros(n + 1, np.int64)
    np.cumsum(np.bincount(rows, minlength=n), out=rp[1:])
    assert np.diff(rp).min() >= 1024
    A = ctx.csr(n, n, rp, col, None)
    want, nlev, edges = orc.bfs_levels(rp, col, None, 11)
    for AT in (A, None):
        levels, st = ctx.bfs(A, 11, AT=AT)
        assert np.array_equal(levels.download(np.int32, n), want) and st.iterations == nlev and st.edges_traversed == edges
    # an operator without entries: only the root is reached
    E = ctx.csr(50, 50, np.zeros(51, np.int64), np.zeros(0, np.int32), None)
    for AT in (E, None):
        levels, st = ctx.bfs(E, 7, AT=AT)
        got = levels.download(np.int32, 50)
        assert got[7] == 0 and (np.delete(got, 7) == -1).all() and st.reached == 1
    dist, st = ctx.sssp(E, 7, AT=E)
    d = dist.download(np.float32, 50)
    assert d[7] == 0 and (np.delete(d, 7) == np.float32(3.4028235e38)).all() and st.converged


def test_sssp_zero_and_equal_weights(ctx, orc):
    """All weights zero (bucket width 0: every bucket holds one distinct distance) and all weights equal."""
    rng = np.random.default_rng(43)
    n = 5000
    rp, col = random_ragged(rng, n, n, mean_len=5, long_rows=[(0, 700)], empty_frac=0.1)
    for w in (0.0, 2.5):
        val = np.full(len(col), w, np.float32)
        A = ctx.csr(n, n, rp, col, val)
        dist, st = ctx.sssp(A, 0)
        assert st.converged
        assert np.array_equal(dist.download(np.float32, n).view(np.int32), orc.sssp(rp, col, val, 0).view(np.int32)), w


def test_sssp_grid_and_directed(ctx, orc):
    G = ctx.synth_grid2d(96, 61)
    rp, col, val = G.download()
    n = len(rp) - 1
    dist, st = ctx.sssp(G, 0, AT=G)  # symmetric weights: G is its own transpose
    assert st.converged and st.reached == n and st.edges_traversed == len(col)
    assert np.array_equal(dist.download(np.float32, n), orc.sssp(rp, col, val, 0))
    # directed graph with non-integer weights: exact as well (monotone rounding)
    rp, col, val = orc.gen_uniform(4000, 4000, 6)
    A = ctx.csr(4000, 4000, rp, col, val)
    dist, st = ctx.sssp(A, 5)
    assert np.array_equal(dist.download(np.float32, 4000).view(np.int32), orc.sssp(rp, col, val, 5).view(np.int32))


def test_sssp_dense_graph_with_few_vertices(ctx, orc, monkeypatch):
    """nnz >> n with weights over five orders of magnitude and ONE huge bucket: every vertex is improved many times in a
    phase, which is what fills the global near list (n + 1024 entries at first; the driver starts over with a list four
    times as long when it overflows instead of failing a valid input).  Exact against the oracle either way."""
    rng = np.random.default_rng(77)
    n, k = 1800, 1500
    rp = np.arange(n + 1, dtype=np.int64) * k
    col = np.concatenate([rng.choice(n, k, replace=False) for _ in range(n)]).astype(np.int32)
    val = np.floor(10.0 ** rng.uniform(0, 5, n * k)).astype(np.float32)
    A = ctx.csr(n, n, rp, col, val)
    want = orc.sssp(rp, col, val, 3)
    for env in ({}, {"SHB_SSSP_DELTA": "1e30"}):
        for key, v in env.items():
            monkeypatch.setenv(key, v)
        dist, st = ctx.sssp(A, 3)
        assert st.converged and np.array_equal(dist.download(np.float32, n).view(np.int32), want.view(np.int32)), env


@pytest.mark.parametrize("env", [{}, {"SHB_SSSP_DELTA": "0.01"}, {"SHB_SSSP_DELTA": "1e30"}, {"SHB_SSSP_C": "2"},
                                 {"SHB_SSSP_CTAS_PER_SM": "2"}])
def test_sssp_near_far_schedules_reach_the_same_fixed_point(ctx, orc, env, monkeypatch):
    """Near-far SSSP on a power-law graph with hub rows (warp-cooperative relaxation, shared-memory list
    spills) and non-integer float weights; any bucket width -- one bucket per distinct distance, a single
    bucket (plain label correcting), the heuristic -- gives the oracle's distances bit for bit."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    A = ctx.synth_rmat(14, 16, seed=orc.SEED_C3, symmetrise=True, value_mode=1)
    rp, col, val = A.download()
    n = len(rp) - 1
    root = int(np.argmax(np.diff(rp) > 0))
    want = orc.sssp(rp, col, val, root)
    dist, st = ctx.sssp(A, root)  # values are not symmetric: the library builds the transpose
    assert st.converged
    assert np.array_equal(dist.download(np.float32, n).view(np.int32), want.view(np.int32))
    assert st.reached == int((want < orc.FLT_MAX).sum())
    # pattern operator: every edge weighs 1 -> distances are BFS levels
    P = ctx.synth_rmat(12, 8, seed=orc.SEED_C3, symmetrise=True)
    prp, pcol, _ = P.download()
    proot = int(np.argmax(np.diff(prp) > 0))
    lv, _, _ = orc.bfs_levels(prp, pcol, None, proot)
    dist, st = ctx.sssp(P, proot, AT=P)
    d = dist.download(np.float32, len(prp) - 1)
    assert st.converged and np.array_equal(d[lv >= 0], lv[lv >= 0].astype(np.float32)) and (d[lv < 0] == orc.FLT_MAX).all()


def test_sssp_phase_cap_reports_not_converged(ctx, orc):
    G = ctx.synth_grid2d(64, 64)
    with pytest.raises(shb.ShbError):
        ctx.sssp(G, 0, AT=G, max_rounds=3)


def test_pagerank_canonical(ctx, orc):
    A = ctx.synth_rmat(12, 16, seed=orc.SEED_C5, symmetrise=False, value_mode=1)
    rp, col, val = A.download()
    n = len(rp) - 1
    want, wi = orc.pagerank(rp, col, val, delta=1e-7, max_iters=300, acc_double=True)
    wf, wif = orc.pagerank(rp, col, val, delta=1e-7, max_iters=300, acc_double=False)
    for cfg in (None, shb.launch_cfg(variant=shb.VARIANT_SCALAR), shb.launch_cfg(variant=shb.VARIANT_STREAM)):
        res, st = ctx.pagerank(A, delta=1e-7, max_iters=300, cfg=cfg)
        got = res.download(np.float32, n)
        assert st.converged and abs(int(st.iterations) - wif) <= 2
        assert np.abs(got - want).max() <= 1e-5 * np.abs(want).max()
        assert np.abs(got - wf).max() <= 1e-5 * np.abs(wf).max()
    # the scalar schedule keeps the CPU loop's order: identical bits and identical sweep count
    res, st = ctx.pagerank(A, delta=1e-7, max_iters=300, cfg=shb.launch_cfg(variant=shb.VARIANT_SCALAR))
    assert st.iterations == wif and np.array_equal(res.download(np.float32, n).view(np.int32), wf.view(np.int32))


def test_scatter_to_several_destinations_row_partitioned(ctx, orc):
    """The multi-GPU kernel on one device: two row blocks with global column ids, each writing its
    slice of the result into BOTH destination vectors (what every rank does over NVLink)."""
    n = 6000
    rp, col, val = orc.gen_uniform(n, n, 16, val_mode=1)
    x = orc.xmode_vector(n, 1)
    want = orc.spmv(orc.PLUS_TIMES, rp, col, val, x, x, 1.0, 0.5)
    xb = ctx.array(x)
    d0, d1 = ctx.empty(n * 4), ctx.empty(n * 4)
    flag = ctx.empty(4); flag.zero()
    cut = 2500
    for lo, hi in ((0, cut), (cut, n)):
        part = ctx.csr(hi - lo, n, rp[lo:hi + 1] - rp[lo], col[rp[lo]:rp[hi]], val[rp[lo]:rp[hi]])
        ctx.spmv_scatter(part, shb.PLUS_TIMES, xb, None, 1.0, 0.5, [d0, d1], lo, delta=1e-4, changed=flag)
    ctx.synchronize()
    assert np.array_equal(d0.download(np.float32, n), want) and np.array_equal(d1.download(np.float32, n), want)
    assert flag.download(np.uint32, 1)[0] == 1
    # fixed point: x = out -> flag stays 0 (identity operator)
    eye = ctx.csr(n, n, np.arange(n + 1), np.arange(n, dtype=np.int32), np.ones(n, np.float32))
    flag.zero()
    ctx.spmv_scatter(eye, shb.PLUS_TIMES, xb, None, 1.0, 0.0, [d0], 0, delta=1e-4, changed=flag)
    ctx.synchronize()
    assert flag.download(np.uint32, 1)[0] == 0 and np.array_equal(d0.download(np.float32, n), x)


@pytest.mark.parametrize("with_transpose", [True, False])
def test_bfs_level_cap_reports_not_converged_with_the_levels_found(ctx, orc, with_transpose):
    """max_levels smaller than the depth of the traversal: SHB_ERR_NOT_CONVERGED (status 7), and the levels reached so
    far are exact (the reference loops forever on such inputs, Q8)."""
    rp, col, _ = orc.gen_grid2d(50, 3)
    n = len(rp) - 1
    A = ctx.csr(n, n, rp, col, None)
    want, nlev, _ = orc.bfs_levels(rp, col, None, 0)
    assert nlev > 20
    with pytest.raises(shb.ShbError) as e:
        ctx.bfs(A, 0, AT=A if with_transpose else None, max_levels=5)
    assert e.value.status == 7
    levels, st = ctx.bfs(A, 0, AT=A if with_transpose else None)
    assert st.converged and st.iterations == nlev and np.array_equal(levels.download(np.int32, n), want)
    # a long traversal laps the driver's count ring (256 frontiers)
    rp2, col2, _ = orc.gen_grid2d(700, 2)
    A2 = ctx.csr(len(rp2) - 1, len(rp2) - 1, rp2, col2, None)
    want2, nlev2, edges2 = orc.bfs_levels(rp2, col2, None, 0)
    lv2, st2 = ctx.bfs(A2, 0, AT=A2 if with_transpose else None)
    assert nlev2 > 600 and st2.iterations == nlev2 and st2.edges_traversed == edges2
    assert np.array_equal(lv2.download(np.int32, len(rp2) - 1), want2)


def test_bfs_pull_mappings_agree(ctx, orc):
    """The compacted pull (default) and the thread-per-row pull (SHB_BFS_PULL=rows, read once per process) give the
    same levels as the queue BFS on an RMAT graph with hubs, empty rows and a ragged last bitmap word; the operator is
    traversed twice so that the cached empty-row bits are used as well as built."""
    import hashlib
    import os
    import subprocess
    import sys

    root_dir = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    A = ctx.synth_rmat(14, 16, seed=orc.SEED_C3, symmetrise=True)
    rp, col, _ = A.download()
    root = int(np.argmax(np.diff(rp) > 0))
    want, nlev, _ = orc.bfs_levels(rp, col, None, root)
    for AT in (A, None, A):
        levels, st = ctx.bfs(A, root, AT=AT)
        assert np.array_equal(levels.download(np.int32, len(rp) - 1), want) and st.iterations == nlev
    script = (
        "import sys, hashlib, numpy as np; sys.path.insert(0, %r); import sparseharness_b200 as shb\n"
        "ctx = shb.Context(0); A = ctx.synth_rmat(14, 16, seed=%d, symmetrise=True)\n"
        "for AT in (A, None):\n"
        "    lv, st = ctx.bfs(A, %d, AT=AT)\n"
        "    print(hashlib.sha256(lv.download(np.int32, 1 << 14).tobytes()).hexdigest(), st.iterations)\n"
    ) % (root_dir, orc.SEED_C3, root)
    r = subprocess.run([sys.executable, "-c", script], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, SHB_BFS_PULL="rows"))
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l.split() for l in r.stdout.strip().splitlines() if len(l.split()) == 2]
    assert len(lines) == 2
    for digest, iters in lines:
        assert digest == hashlib.sha256(want.astype(np.int32).tobytes()).hexdigest() and int(iters) == nlev


def test_power_iteration_matches_oracle(ctx, orc, mtx_dir):
    """Canonical eigenvector driver (shb_power_iteration): SpMV, infinity norm, fused scale + convergence test, all on
    the device.  SCALAR schedule == the oracle's loop bit for bit (same adds, same IEEE divide); the default schedule
    to 1e-5; the eigenvalue estimate is the oracle's."""
    rp, col, val = orc.gen_uniform(5000, 5000, 12, val_mode=1)
    val = (np.abs(val) + 1).astype(np.float32)
    n = len(rp) - 1
    A = ctx.csr(n, n, rp, col, val)
    x0 = np.ones(n, np.float32)
    want, wev, wit = orc.power_iteration(rp, col, val, x0, delta=1e-5, max_iters=100)
    got, ev, st = ctx.power_iteration(A, x0, delta=1e-5, max_iters=100, cfg=shb.launch_cfg(variant=shb.VARIANT_SCALAR))
    assert st.converged and st.iterations == wit and ev == np.float32(wev)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    got, ev, st = ctx.power_iteration(A, x0, delta=1e-5, max_iters=100)
    assert st.converged and abs(int(st.iterations) - wit) <= 1 and abs(ev - wev) <= 1e-5 * wev
    assert np.abs(got - want).max() <= 1e-4
    # the cap: 3 iterations, not converged, still the oracle's vector
    want3, _, wit3 = orc.power_iteration(rp, col, val, x0, delta=1e-9, max_iters=3)
    got3, _, st3 = ctx.power_iteration(A, x0, delta=1e-9, max_iters=3, cfg=shb.launch_cfg(variant=shb.VARIANT_SCALAR))
    assert wit3 == 3 and st3.iterations == 3 and not st3.converged and np.array_equal(got3.view(np.uint32), want3.view(np.uint32))
    # a fixture through the reference's loader (integer-valued): every schedule bit for bit
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir["Trefethen_20"], "float")
    A = ctx.csr(n, n, rp, col, val)
    want, wev, wit = orc.power_iteration(rp, col, val, np.ones(n, np.float32), delta=1e-4, max_iters=10)
    got, ev, st = ctx.power_iteration(A, np.ones(n, np.float32), delta=1e-4, max_iters=10, cfg=shb.launch_cfg(variant=shb.VARIANT_SCALAR))
    assert st.iterations == wit and np.array_equal(got.view(np.uint32), want.view(np.uint32)) and ev == np.float32(wev)
