"""-m gpu: BASELINE.json's configurations at their stated sizes.

config 2 (16M x 16 SpMV): sampled row blocks against the oracle + agreement of independent schedules on all rows.
config 3 (RMAT-24 BFS) and config 4 (8192^2 grid SSSP): the WHOLE operator is downloaded and the oracle's queue BFS /
Dijkstra run on it (CPU minutes, once per suite); all 16.7M levels / 67.1M distances of the single-device driver AND
of the row-partitioned drivers (kernels and protocol of the multi-GPU path, blocks emulated on this device) must
equal the oracle's bit for bit.
config 5 (RMAT-26 PageRank): after the 20 sweeps, x_20 on sampled row blocks is re-computed from x_19 by the
oracle's plus-times loop (1e-5 relative), single device and inside the row-partitioned run."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import numpy as np
import pytest

import sparseharness_b200 as shb

pytestmark = pytest.mark.gpu
N2 = 1 << 24  # config 2: 16M rows x 16 nnz/row


@pytest.fixture(scope="module")
def ctx():
    c = shb.Context(0)
    yield c
    c.close()


def test_config2_full_size_spmv(ctx, orc):
    out = ctx.empty(N2 * 4)
    for val_mode, xmode in ((1, 1), (0, 0)):  # integer-valued (exact) and uniform[-1,1) floats (1e-5)
        A = ctx.synth_uniform(N2, N2, 16, val_mode=val_mode)
        info = A.info
        assert info.nnz == N2 * 16 and info.algorithmic_bytes_spmv == 2348810244 and info.max_row_len == 16
        x = ctx.synth_vector(N2, mode=xmode)
        xh = x.download(np.float32, N2)
        results = {}
        geom = A.build_blocked(make_default=False)  # config 2 at full size: 4 column passes of 148 slices
        assert geom["n_pass"] == 4 and geom["W"] <= 30720 and geom["entries"] >= info.nnz
        for label, cfg in (("auto", {}), ("scalar", dict(variant=shb.VARIANT_SCALAR)),
                           ("stream", dict(variant=shb.VARIANT_STREAM)), ("vector8", dict(variant=shb.VARIANT_VECTOR, lanes=8)),
                           ("blocked", dict(variant=shb.VARIANT_BLOCKED))):
            out.fill_u32(0xDEADBEEF)
            ctx.spmv(A, shb.PLUS_TIMES, x, None, 1.0, 0.0, out, shb.launch_cfg(**cfg))
            results[label] = out.download(np.float32, N2)
        # sampled row blocks against the oracle (first, middle, last)
        for begin in (0, N2 // 2 - 1234, N2 - 4096):
            rp, col, val = A.download(begin, 4096)
            orp, ocol, oval = orc.gen_uniform(N2, N2, 16, val_mode=val_mode, row_begin=begin, row_count=4096)
            assert np.array_equal(col, ocol) and np.array_equal(val.view(np.int32), oval.view(np.int32))
            want = orc.spmv(orc.PLUS_TIMES, rp, col, val, xh, np.zeros(4096, np.float32), 1.0, 0.0)
            for label, r in results.items():
                got = r[begin:begin + 4096]
                if val_mode == 1 or label in ("scalar", "stream"):
                    assert np.array_equal(got, want), (label, begin)
                else:
                    assert np.abs(got - want).max() <= 1e-5 * np.abs(want).max(), (label, begin)
        # independent schedules agree on ALL 16M rows
        assert np.array_equal(results["scalar"], results["stream"])
        if val_mode == 1:
            assert np.array_equal(results["auto"], results["scalar"]) and np.array_equal(results["vector8"], results["scalar"])
            assert np.array_equal(results["blocked"], results["scalar"])
            assert results["scalar"].astype(np.float64).sum() == results["auto"].astype(np.float64).sum()
        else:
            s = np.abs(results["scalar"]).max()
            assert np.abs(results["auto"] - results["scalar"]).max() <= 1e-5 * s
            assert np.abs(results["vector8"] - results["scalar"]).max() <= 1e-5 * s
            assert np.abs(results["blocked"] - results["scalar"]).max() <= 1e-5 * s
        A.free(); x.free()


def test_config4_grid_sssp_against_dijkstra_4m_vertices(ctx, orc):
    """2048 x 2048 slice of config 4's grid family: frontier SSSP == CPU Dijkstra bit for bit, and
    distances satisfy the Bellman optimality conditions (size-independent property)."""
    W = H = 2048
    G = ctx.synth_grid2d(W, H)
    dist, st = ctx.sssp(G, 0, AT=G)
    n = W * H
    d = dist.download(np.float32, n)
    rp, col, val = G.download()
    assert st.converged and st.reached == n and st.edges_traversed == len(col)
    assert np.array_equal(d, orc.sssp(rp, col, val, 0))
    # optimality: d[r] == min_k d[col k] + w_k for r != source, and no edge can relax further
    cand = d[col] + val
    rowmin = np.minimum.reduceat(cand, rp[:-1])
    assert d[0] == 0 and np.array_equal(rowmin[1:], d[1:])


def test_config3_bfs_scale20_properties(ctx, orc):
    """RMAT scale 20 (config 3's family): levels == CPU queue BFS; every edge spans at most one level."""
    A = ctx.synth_rmat(20, 16, seed=orc.SEED_C3, symmetrise=True)
    rp, col, _ = A.download()
    n = len(rp) - 1
    root = int(np.argmax(np.diff(rp) > 0))
    levels, st = ctx.bfs(A, root, AT=A)
    lv = levels.download(np.int32, n)
    want, nlev, edges = orc.bfs_levels(rp, col, None, root)
    assert np.array_equal(lv, want) and st.iterations == nlev and st.edges_traversed == edges
    rows = np.repeat(np.arange(n), np.diff(rp))
    reached = lv[rows] >= 0
    assert (lv[col][reached] >= 0).all() and np.abs(lv[rows][reached] - lv[col][reached]).max() <= 1


def test_config3_full_size_bfs_equals_oracle_on_every_vertex(ctx, orc):
    """RMAT scale 24 (config 3 at BASELINE's size): the oracle's queue BFS on the downloaded operator gives all
    16.7M levels; shb_bfs and the row-partitioned BFS on 4 blocks (emulated on this device) must both equal them."""
    from sparseharness_b200 import multi
    A = ctx.synth_rmat(24, 16, seed=0x5EED0024, symmetrise=True)
    info = A.info
    n = info.n_rows
    rp_head = A.download(0, 4096)[0]
    root = int(np.argmax(np.diff(rp_head) > 0))
    levels, st = ctx.bfs(A, root, AT=A)
    single = levels.download(np.int32, n)
    levels.free()
    rp_all, col_all, _ = A.download()
    assert len(col_all) == info.nnz
    want, nlev, edges = orc.bfs_levels(rp_all, col_all, None, root)  # the oracle, every vertex
    del rp_all, col_all
    assert np.array_equal(single, want) and st.iterations == nlev and st.edges_traversed == edges
    assert st.reached == int((want >= 0).sum())
    deg = A.row_lengths()
    bounds = multi.word_aligned_bounds(A.balance_rows(4))
    blocks = [A.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(4)]
    be = multi.EmulatedBfsPartition(ctx, blocks, bounds, root, info.nnz, at_blocks=blocks, deg=deg)
    n_levels, _, _ = multi.run_bfs(be, 1 << 20, 8)
    got = be.result()
    assert n_levels == st.iterations and np.array_equal(got, want)
    assert be.stats() == (st.reached, st.edges_traversed)
    # size-independent property on a row sample: neighbours of a reached vertex are reached, one level apart at most
    rp, col, _ = A.download(n // 2, 1 << 16)
    rows = np.repeat(np.arange(n // 2, n // 2 + (1 << 16)), np.diff(rp))
    r = got[rows] >= 0
    assert (got[col][r] >= 0).all() and (np.abs(got[rows][r] - got[col][r]) <= 1).all()


def test_config4_full_size_sssp_equals_oracle_on_every_vertex(ctx, orc):
    """8192 x 8192 grid (config 4 at BASELINE's size): the oracle's Dijkstra (same float adds) on the downloaded
    operator gives all 67.1M distances; shb_sssp and the block-asynchronous near-far on 2 blocks with distance
    windows (emulated on this device) must both equal them bit for bit."""
    from sparseharness_b200 import multi
    G = ctx.synth_grid2d(8192, 8192)
    n = G.info.n_rows
    dist, st = ctx.sssp(G, 0, AT=G)
    single = dist.download(np.float32, n)
    dist.free()
    rp_all, col_all, val_all = G.download()
    want = orc.sssp(rp_all, col_all, val_all, 0)  # the oracle, every vertex
    del rp_all, col_all, val_all
    assert st.converged and np.array_equal(single.view(np.uint32), want.view(np.uint32))
    assert st.reached == int((want < orc.FLT_MAX).sum())
    bounds = G.balance_rows(2)
    blocks = [G.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(2)]
    be = multi.EmulatedSsspPartition(ctx, blocks, bounds, 0, window_buckets=8.0)
    rounds, conv = multi.run_sssp(be, 4096, 8)
    got = be.result()
    assert conv and rounds > 2 and np.array_equal(got, want)
    assert be.stats_total() == (st.reached, st.edges_traversed)
    lo = int(bounds[1]) - 4096  # rows either side of the block boundary
    rp, col, val = G.download(lo, 8192)
    cand = got[col] + val
    assert np.array_equal(np.minimum.reduceat(cand, rp[:-1]), got[lo:lo + 8192])


def test_config5_full_size_pagerank_sweep_against_oracle(ctx, orc):
    """RMAT scale 26, 2^30 non-zeros (config 5 at BASELINE's size), 20 sweeps on one device and on a 2-block row
    partition (emulated): x_20 on sampled row blocks == the oracle's plus-times loop applied to x_19 (1e-5 relative),
    x_20 of the two runs agree to 1e-5, and x_20 sums to ~1 minus the mass lost in dangling columns."""
    import parity_checks as pc
    from sparseharness_b200 import multi
    A = ctx.synth_rmat(26, 16, seed=0x5EED0026, symmetrise=False, value_mode=1)
    info = A.info
    n = info.n_rows
    assert n == 1 << 26 and info.nnz > (1 << 30) * 0.9
    beta = float((np.float32(1.0) - np.float32(0.85)) / np.float32(n))
    res, prev, st = ctx.pagerank(A, delta=0.0, max_iters=20, check_every=20, both=True)
    x20, x19 = res.download(np.float32, n), prev.download(np.float32, n)
    res.free(); prev.free()
    assert st.iterations == 20
    begins, cnt = pc.sample_blocks(n, 4096, extra=[n // 3, 12345])
    for b in begins:
        rp, col, val = A.download(b, cnt)
        assert pc.pagerank_rows_err(orc, x19, x20[b:b + cnt], rp, col, val, beta) <= 1e-5, b
    assert np.isfinite(x20).all() and 0.1 < float(x20.astype(np.float64).sum()) <= 1.0 + 1e-3
    # the row-partitioned recurrence (2 blocks emulated on this device), same 20 sweeps
    bounds = A.balance_rows(2)
    blocks = [A.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(2)]
    A.free()
    be = multi.EmulatedRowPartition(ctx, blocks, bounds, shb.PLUS_TIMES, 1.0, beta, y_mode="one", delta=0.0, term=shb.TERM_ABSDIFF,
                                    max_iters=32)
    be.x[0].fill_u32(int(np.array([np.float32(1.0) / np.float32(n)], np.float32).view(np.uint32)[0]), n)
    iters, conv, idx = multi.run_fixed_point(be, 20, 20)
    assert iters == 20
    p20, p19 = be.result(idx), be.result(idx ^ 1)
    assert np.abs(p20 - x20).max() <= 1e-5 * np.abs(x20).max()
    for k in range(2):
        lo = int(bounds[k])
        rp, col, val = blocks[k].download(100, 4096)
        assert pc.pagerank_rows_err(orc, p19, p20[lo + 100:lo + 100 + 4096], rp, col, val, beta) <= 1e-5, k
