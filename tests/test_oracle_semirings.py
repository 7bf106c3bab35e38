"""Internal consistency of the CPU oracle on the semirings the reference has no CPU code for
(min-plus, or-and, max-min: 'parity unpinned by the reference'), plus the generators."""
import numpy as np
import pytest


def test_sssp_faithful_iteration_equals_dijkstra_on_fixture(orc, mtx_dir):
    # scripts/example/sssp.sh runs sssp on matrix3 (Trefethen_20)
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir["Trefethen_20"], "float")
    x0 = np.full(n, orc.FLT_MAX, np.float32); x0[0] = 0
    res, iters = orc.iterate(orc.MIN_PLUS, rp, col, val, x0, x0.copy(), 0.0, 0.0, delta=1e-4)
    assert 2 <= iters < 30
    assert np.array_equal(res, orc.sssp(rp, col, val, 0))
    # both accumulation orders give the same bits (min is order independent)
    res2, _ = orc.iterate(orc.MIN_PLUS, rp, col, val, x0, x0.copy(), 0.0, 0.0, delta=1e-4, order=1)
    assert np.array_equal(res, res2)


def test_sssp_grid_bellman_ford_equals_dijkstra(orc):
    rp, col, val = orc.gen_grid2d(24, 17)
    n = len(rp) - 1
    x0 = np.full(n, orc.FLT_MAX, np.float32); x0[0] = 0
    res, iters = orc.iterate(orc.MIN_PLUS, rp, col, val, x0, x0.copy(), 0.0, 0.0, delta=1e-4, threads=4)
    d = orc.sssp(rp, col, val, 0)
    assert np.array_equal(res, d) and iters >= 24 + 17 - 2
    assert d.max() < orc.FLT_MAX and (val >= 1).all() and (val <= 255).all()
    # symmetric weights: w(u,v) == w(v,u)
    w = {}
    for r in range(n):
        for k in range(rp[r], rp[r + 1]):
            w[(r, col[k])] = val[k]
    assert all(w[(b, a)] == v for (a, b), v in w.items())


def test_bfs_levels_match_frontier_recurrence(orc):
    rp, col = orc.gen_rmat(10, 8, symmetrise=True)
    n = len(rp) - 1
    assert (np.diff(rp) >= 0).all()
    deg = np.diff(rp)
    root = int(np.argmax(deg > 0))
    level, nlev, edges = orc.bfs_levels(rp, col, None, root)
    # frontier_{k+1} = (A (x) frontier_k) AND NOT visited, with the or-and kernel restatement
    ones = np.ones(len(col), np.int32)
    front = np.zeros(n, np.int32); front[root] = 1
    lev = np.full(n, -1, np.int32); lev[root] = 0
    k = 0
    while front.any():
        nxt = orc.spmv(orc.OR_AND, rp, col, ones, front, front, 1, 0)
        nxt = np.where(lev >= 0, 0, nxt).astype(np.int32)
        k += 1
        lev[nxt == 1] = k
        front = nxt
    assert np.array_equal(lev, level) and nlev == k
    assert edges == int(deg[level >= 0].sum())
    # symmetrised + de-duplicated + no self loops
    for r in range(0, n, 37):
        c = col[rp[r]:rp[r + 1]]
        assert (np.diff(c) > 0).all() and (c != r).all()


def test_pagerank_float_vs_double_accumulation(orc):
    rp, col = orc.gen_rmat(10, 16, seed=orc.SEED_C5, symmetrise=False)
    val = orc.pagerank_values(rp, col)
    a, ia = orc.pagerank(rp, col, val, delta=1e-7, max_iters=200, acc_double=False)
    b, ib = orc.pagerank(rp, col, val, delta=1e-7, max_iters=200, acc_double=True)
    assert ia < 200 and ib < 200
    assert np.abs(a - b).max() <= 1e-5 * np.abs(b).max()
    # column-stochastic * d: every column with entries sums to 0.85
    cs = np.zeros(len(rp) - 1); np.add.at(cs, col, val.astype(np.float64))
    nz = cs > 0
    assert np.allclose(cs[nz], 0.85, rtol=1e-5)


def test_scc_faithful_iteration_terminates_on_fixture(orc, mtx_dir):
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir["Trefethen_20"], "int", normalise="scc")
    x0 = np.arange(n, dtype=np.int32)
    y0 = np.full(n, orc.INT_MIN, np.int32)
    res, iters = orc.iterate(orc.MAX_MIN, rp, col, val, x0, y0, orc.INT_MAX, orc.INT_MIN, max_iters=100)
    assert iters < 100
    again = orc.spmv(orc.MAX_MIN, rp, col, val, res, res, orc.INT_MAX, orc.INT_MIN)
    assert np.array_equal(again, res)  # a fixed point


def test_uniform_generator_shape_and_determinism(orc):
    rp, col, val = orc.gen_uniform(1000, 1000, 16)
    assert np.array_equal(rp, np.arange(1001) * 16)
    c = col.reshape(1000, 16)
    assert (np.diff(c, axis=1) >= 0).all() and c.min() >= 0 and c.max() < 1000
    assert val.min() >= -1 and val.max() < 1
    rp2, col2, val2 = orc.gen_uniform(1000, 1000, 16, row_begin=100, row_count=50)
    assert np.array_equal(col2, col[1600:2400]) and np.array_equal(val2, val[1600:2400])
    _, _, vi = orc.gen_uniform(1000, 1000, 16, val_mode=1)
    assert set(np.unique(vi)) <= set(range(-8, 9))


def test_gold_beta_quirk_is_reproduced_only_for_beta_zero(orc, mtx_dir):
    """Q5: gold's beta term is per non-zero; the device formula is per row.  They agree for beta = 0."""
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir["GD96_b"], "float")
    x = orc.xmode_vector(n, 1)
    y = np.full(n, 2.0, np.float32)
    assert np.array_equal(orc.gold_spmv(rp, col, val, x, y, 1.0, 0.0), orc.spmv(orc.PLUS_TIMES, rp, col, val, x, y, 1.0, 0.0))
    assert not np.array_equal(orc.gold_spmv(rp, col, val, x, y, 1.0, 0.5), orc.spmv(orc.PLUS_TIMES, rp, col, val, x, y, 1.0, 0.5))
