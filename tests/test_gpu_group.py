"""-m gpu: the multi-GPU drivers INSIDE the boundary (shb_group_*, include/shb.h): one host thread, N devices, no
torch.distributed.  A group of one device runs everywhere (it exercises the whole loop: scatter launch, publish
kernel, host polls); the tests that need a second device skip on a one-GPU box.  Oracle = the checker."""
import numpy as np
import pytest

import sparseharness_b200 as shb

pytestmark = pytest.mark.gpu


def _n_dev():
    return shb.device_count()


def _sizes():
    return [1] + ([2] if _n_dev() >= 2 else []) + ([4] if _n_dev() >= 4 else [])


@pytest.mark.parametrize("n_dev", [1, 2, 4])
def test_group_pagerank_bfs_sssp_match_oracle(orc, n_dev):
    if _n_dev() < n_dev:
        pytest.skip(f"needs {n_dev} GPUs")
    g = shb.Group(n_dev)
    # PageRank, directed RMAT with d / colsum values (canonical mode)
    rp, col = orc.gen_rmat(12, 8, seed=orc.SEED_C3, symmetrise=False)
    val = orc.pagerank_values(rp, col)
    n = len(rp) - 1
    part = g.partition_host(n, n, rp, col, val, symmetric=False)
    b = part.info["bounds"]
    assert b[0] == 0 and b[-1] == n and (np.diff(b) >= 0).all() and all(int(v) % 32 == 0 for v in b[1:-1])
    x20, x19, st = g.pagerank(part, delta=0.0, max_iters=20, check_every=7, want_prev=True)
    want20, _ = orc.pagerank(rp, col, val, delta=0.0, max_iters=20, acc_double=True)
    want19, _ = orc.pagerank(rp, col, val, delta=0.0, max_iters=19, acc_double=True)
    assert st.iterations == 20 and not st.converged
    assert np.abs(x20 - want20).max() <= 1e-5 * np.abs(want20).max()
    assert np.abs(x19 - want19).max() <= 1e-5 * np.abs(want19).max()
    res, _, st = g.pagerank(part, delta=1e-7, max_iters=400, check_every=5)
    want, wit = orc.pagerank(rp, col, val, delta=1e-7, max_iters=400)
    assert st.converged and abs(int(st.iterations) - wit) <= 1 and np.abs(res - want).max() <= 1e-5 * np.abs(want).max()
    part.free()
    # BFS levels, symmetrised RMAT
    rp, col = orc.gen_rmat(13, 8, seed=orc.SEED_C3, symmetrise=True)
    n = len(rp) - 1
    root = int(np.argmax(np.diff(rp) > 0))
    part = g.partition_host(n, n, rp, col, None, symmetric=True)
    lv, st = g.bfs(part, root, check_every=3)
    want, nlev, edges = orc.bfs_levels(rp, col, None, root)
    assert np.array_equal(lv, want) and st.iterations == nlev and st.edges_traversed == edges and st.reached == int((want >= 0).sum())
    part.free()
    # BFS on a DIRECTED graph: the partition builds the out-edge blocks itself
    rp, col = orc.gen_rmat(11, 8, seed=orc.SEED_C3, symmetrise=False)
    n = len(rp) - 1
    deg_out = np.bincount(col, minlength=n)
    root = int(np.argmax(deg_out > 0))
    part = g.partition_host(n, n, rp, col, None, symmetric=False)
    lv, st = g.bfs(part, root)
    want, nlev, _ = orc.bfs_levels(rp, col, None, root)
    assert np.array_equal(lv, want) and st.iterations == nlev
    part.free()
    # SSSP on a grid (high diameter): exact distances
    rp, col, val = orc.gen_grid2d(96, 80)
    n = len(rp) - 1
    part = g.partition_host(n, n, rp, col, val, symmetric=True)
    d, st = g.sssp(part, 0, check_every=4)
    want = orc.sssp(rp, col, val, 0)
    assert st.converged and np.array_equal(d.view(np.uint32), want.view(np.uint32)) and st.reached == n
    d2, st2 = g.sssp(part, 0, window_buckets=0.0)
    assert st2.converged and np.array_equal(d2.view(np.uint32), want.view(np.uint32))
    part.free()
    g.close()


@pytest.mark.parametrize("n_dev", [1, 2])
def test_group_iterate_faithful_recurrences(orc, mtx_dir, n_dev):
    """The reference's executeRun recurrence (app/sssp.cpp:112-153, y re-bound to x) on the partition: same launch
    count and same bits as the oracle, which is pinned to the reference's own kernel text (tests/test_oracle_pinned.py)."""
    if _n_dev() < n_dev:
        pytest.skip(f"needs {n_dev} GPUs")
    g = shb.Group(n_dev)
    for name in ("1138_bus", "Trefethen_20", "GD96_b"):
        n, _, rp, col, val = orc.mm_load_csr(mtx_dir[name], "float")
        x0 = np.full(n, orc.FLT_MAX, np.float32); x0[0] = 0
        part = g.partition_host(n, n, rp, col, val)
        res, _, it, ns, ok = g.iterate(part, shb.MIN_PLUS, x0, x0.copy(), 0.0, 0.0, delta=1e-4, max_iters=400, check_every=5)
        want, wit = orc.iterate(orc.MIN_PLUS, rp, col, val, x0, x0.copy(), 0.0, 0.0, delta=1e-4, max_iters=400)
        assert ok and it == wit and ns > 0 and np.array_equal(res.view(np.uint32), want.view(np.uint32)), name
        part.free()
        n, _, rp, col, val = orc.mm_load_csr(mtx_dir[name], "int", normalise="scc")
        x0 = np.arange(n, dtype=np.int32); y0 = np.full(n, orc.INT_MIN, np.int32)
        part = g.partition_host(n, n, rp, col, val)
        res, _, it, _, ok = g.iterate(part, shb.MAX_MIN, x0, y0, int(orc.INT_MAX), int(orc.INT_MIN), max_iters=40)
        want, wit = orc.iterate(orc.MAX_MIN, rp, col, val, x0, y0, orc.INT_MAX, orc.INT_MIN, max_iters=40)
        assert it == wit and np.array_equal(res, want), name
        part.free()
    g.close()


def test_group_adopts_device_built_blocks(orc):
    """Blocks generated on the group's own devices (the route the full-size configs take): every device builds the
    graph, keeps its slice; results equal the single-device drivers'."""
    n_dev = min(2, _n_dev())
    g = shb.Group(n_dev)
    blocks, bounds, root = [], None, 0
    for i in range(n_dev):
        c = g.ctx(i)
        A = c.synth_rmat(14, 8, seed=orc.SEED_C3, symmetrise=True)
        if bounds is None:
            from sparseharness_b200 import multi
            bounds = multi.word_aligned_bounds(A.balance_rows(n_dev))
            rp0 = A.download(0, 4096)[0]
            root = int(np.argmax(np.diff(rp0) > 0))
            ref_levels, ref_st = c.bfs(A, root, AT=A)
            want = ref_levels.download(np.int32, A.info.n_rows)
        blocks.append(A.slice_rows(int(bounds[i]), int(bounds[i + 1] - bounds[i])))
        n = A.info.n_rows
        A.free()
    part = g.adopt(blocks, bounds, n)
    lv, st = g.bfs(part, root)
    assert np.array_equal(lv, want) and st.iterations == ref_st.iterations and st.edges_traversed == ref_st.edges_traversed
    part.free()
    g.close()


@pytest.mark.parametrize("n_dev", [1, 2])
def test_group_bfs_level_cap_labels_the_last_frontier(orc, n_dev):
    """max_levels below the depth of the traversal: SHB_ERR_NOT_CONVERGED, and every level up to the cap is exact -- the
    last frontier built (by a push level, which leaves the labelling to the next level's prep) included, as in shb_bfs."""
    import ctypes as C

    from sparseharness_b200._capi import GraphStats, lib
    if _n_dev() < n_dev:
        pytest.skip(f"needs {n_dev} GPUs")
    rp, col, _ = orc.gen_grid2d(60, 40)
    n = len(rp) - 1
    want, nlev, _ = orc.bfs_levels(rp, col, None, 0)
    g = shb.Group(n_dev)
    part = g.partition_host(n, n, rp, col, None, symmetric=True)
    for cap in (1, 5, 12):
        lv = np.full(n, -7, np.int32)
        st = GraphStats()
        status = lib.shb_group_bfs(g.h, part.h, 0, lv.ctypes.data_as(C.c_void_p), cap, 4, C.byref(st))
        assert status == 7 and cap < nlev  # SHB_ERR_NOT_CONVERGED
        assert np.array_equal(lv, np.where(want <= cap, want, -1)), cap
    lv, st = g.bfs(part, 0)
    assert st.converged and np.array_equal(lv, want)
    part.free()
    g.close()
