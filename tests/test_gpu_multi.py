"""GPU tests of the row-partitioned iteration (sparseharness_b200/multi.py): all row blocks emulated on one
device (always), and the real one-process-per-GPU path over NCCL / CUDA IPC when the box has >= 2 GPUs."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import sparseharness_b200 as shb
    c = shb.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("parts", [1, 2, 5])
def test_emulated_row_partition_pagerank_matches_oracle(ctx, orc, parts):
    import sparseharness_b200 as shb
    from sparseharness_b200 import multi
    rp, col = orc.gen_rmat(12, 8, seed=orc.SEED_C5, symmetrise=False)
    n = len(rp) - 1
    val = orc.pagerank_values(rp, col)
    A = ctx.csr(n, n, rp, col, val)
    bounds = A.balance_rows(parts)
    assert np.array_equal(bounds, multi.balanced_bounds(rp, parts))
    blocks = [A.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(parts)]
    for k, B in enumerate(blocks):  # the slice is the same operator rows, rebased, global column ids
        lo, hi = int(bounds[k]), int(bounds[k + 1])
        brp, bcol, bval = B.download()
        assert np.array_equal(brp, rp[lo:hi + 1] - rp[lo]) and np.array_equal(bcol, col[rp[lo]:rp[hi]])
        assert np.array_equal(bval, val[rp[lo]:rp[hi]])
    beta = (np.float32(1.0) - np.float32(0.85)) / np.float32(n)
    be = multi.EmulatedRowPartition(ctx, blocks, bounds, shb.PLUS_TIMES, 1.0, float(beta), y_mode="one", delta=1e-7,
                                    term=shb.TERM_ABSDIFF, cfg=shb.launch_cfg(variant=shb.VARIANT_SCALAR), max_iters=400)
    be.x[0].fill_u32(int(np.array([np.float32(1.0) / np.float32(n)], np.float32).view(np.uint32)[0]), n)
    iters, conv, idx = multi.run_fixed_point(be, 400, 7)
    want, wit = orc.pagerank(rp, col, val, delta=1e-7, max_iters=400)
    assert conv and iters == wit
    assert np.array_equal(be.result(idx).view(np.int32), want.view(np.int32))  # scalar schedule = CPU loop order


def test_emulated_row_partition_faithful_min_plus(ctx, orc):
    """The reference's sssp recurrence (y re-bound to x) on 3 row blocks == the single-process oracle."""
    import sparseharness_b200 as shb
    from sparseharness_b200 import multi
    rp, col, val = orc.gen_grid2d(40, 30)
    n = len(rp) - 1
    A = ctx.csr(n, n, rp, col, val)
    bounds = A.balance_rows(3)
    blocks = [A.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(3)]
    x0 = np.where(np.arange(n) == 0, 0, orc.FLT_MAX).astype(np.float32)
    y0 = ctx.array(x0)
    be = multi.EmulatedRowPartition(ctx, blocks, bounds, shb.MIN_PLUS, 0.0, 0.0, y_mode="rebind", y0=y0, delta=1e-4,
                                    term=shb.TERM_ABSDIFF, max_iters=400)
    be.x[0].upload(x0)
    iters, conv, idx = multi.run_fixed_point(be, 400, 16)
    want, launches = orc.iterate(orc.MIN_PLUS, rp, col, val, x0, x0, 0.0, 0.0, delta=1e-4)
    assert conv and iters == launches and np.array_equal(be.result(idx), want)
    assert np.array_equal(want, orc.sssp(rp, col, val, 0))


@pytest.mark.parametrize("mode", ["nccl", "fused"])
def test_two_gpu_pagerank_matches_oracle(mode):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 29600 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tools", "multi_pagerank.py"), "--scale", "16",
           "--iters", "200", "--delta", "1e-7", "--mode", mode, "--check", "--check-every", "8"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert out["parity_ok"] and out["converged"] and out["n_gpus"] == 2
    assert abs(out["iterations"] - out["oracle_iterations"]) <= 1


@pytest.mark.parametrize("parts,check_every,push,publish", [(1, 4, True, True), (2, 1, True, True), (5, 3, True, False),
                                                            (3, 2, False, True), (8, 8, True, True)])
def test_emulated_row_partition_bfs_matches_queue_bfs(ctx, orc, parts, check_every, push, publish):
    """shb_bfs_block_level on P row blocks of one device (own bitmaps / state per block, stores into every block's
    next bitmap): levels, level count, reached and the TEPS numerator equal the CPU queue BFS, with push levels
    (direction-optimising) and pull-only.  Hub rows >= 128 entries take the warp-cooperative pull path."""
    from sparseharness_b200 import multi
    rp, col = orc.gen_rmat(13, 16, symmetrise=True)
    n = len(rp) - 1
    assert np.diff(rp).max() >= 128
    root = int(np.argmax(np.diff(rp) > 0))
    A = ctx.csr(n, n, rp, col, None)
    deg = A.row_lengths()
    assert np.array_equal(deg.download(np.int32, n), np.diff(rp).astype(np.int32))
    bounds = multi.word_aligned_bounds(A.balance_rows(parts))
    blocks = [A.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(parts)]
    be = multi.EmulatedBfsPartition(ctx, blocks, bounds, root, len(col), at_blocks=blocks if push else None, deg=deg,
                                    publish=publish)
    n_levels, _, launches = multi.run_bfs(be, 1 << 20, check_every)
    want, nlev, edges = orc.bfs_levels(rp, col, None, root)
    assert np.array_equal(be.result(), want)
    assert n_levels == nlev and be.stats() == (int((want >= 0).sum()), edges)
    assert launches == nlev and be.launches >= launches * parts
    # same answer as the single-device direction-optimising driver
    lv, st = ctx.bfs(A, root, AT=A)
    assert np.array_equal(lv.download(np.int32, n), want) and st.edges_traversed == edges
    # a second traversal after reset gives the same levels
    be.reset()
    assert multi.run_bfs(be, 1 << 20, 3)[0] == nlev and np.array_equal(be.result(), want)


@pytest.mark.parametrize("push", [True, False])
def test_emulated_row_partition_bfs_directed_zero_valued_edges_and_level_cap(ctx, orc, push):
    from sparseharness_b200 import multi
    rng = np.random.default_rng(5)
    n = 1000
    rp, col, _ = orc.gen_uniform(n, n, 3, val_mode=1)
    val = rng.integers(0, 2, len(col)).astype(np.int32)  # a stored 0 is no edge (bool_and)
    A = ctx.csr(n, n, rp, col, val)
    AT = A.transpose()
    deg = AT.row_lengths()
    bounds = multi.word_aligned_bounds(A.balance_rows(3))
    blocks = [A.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(3)]
    at_blocks = [AT.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(3)] if push else None
    be = multi.EmulatedBfsPartition(ctx, blocks, bounds, 3, len(col), at_blocks=at_blocks, deg=deg if push else None)
    n_levels, _, _ = multi.run_bfs(be, 1 << 20, 2)
    want, nlev, edges = orc.bfs_levels(rp, col, val, 3)
    assert np.array_equal(be.result(), want) and n_levels == nlev and be.stats()[1] == edges
    # a level cap stops early with every level up to the cap exact: the frontier the last launch built included (a push
    # level leaves its labelling to the next level's prep; the driver labels it when the cap cuts that off)
    be.reset()
    n_levels, _, launches = multi.run_bfs(be, 2, 2)
    got = be.result()
    assert launches == 2 and np.array_equal(got, np.where((want >= 0) & (want <= 2), want, -1))


def test_bfs_block_rejects_unaligned_blocks(ctx, orc):
    import sparseharness_b200 as shb
    rp, col = orc.gen_rmat(8, 4, symmetrise=True)
    n = len(rp) - 1
    A = ctx.csr(n, n, rp, col, None)
    B = A.slice_rows(16, 64)
    nb = n // 8 + 4
    lv, b0, b1, b2, vis, cnt, st = (ctx.empty(64 * 4), ctx.empty(nb), ctx.empty(nb), ctx.empty(nb), ctx.empty(nb),
                                    ctx.empty(8192), ctx.empty(64))
    with pytest.raises(shb.ShbError) as e:
        ctx.bfs_block_init(B, None, 16, n, len(col), 0, None, lv, b0, b1, b2, vis, cnt, st)
    assert e.value.status == 1


@pytest.mark.parametrize("exchange", ["peer", "nccl"])
def test_two_gpu_bfs_matches_oracle(exchange):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 29900 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tools", "multi_bfs.py"), "--scale", "16", "--check", "--exchange", exchange]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])


def test_handles_may_outlive_their_context(ctx, orc):
    """Buffers / operators freed after their context was destroyed (garbage-collected callers free in any order)
    must not touch the dead context nor leave a CUDA error behind for the next call."""
    import sparseharness_b200 as shb
    c2 = shb.Context(0)
    rp, col, val = orc.gen_uniform(256, 256, 4, val_mode=1)
    A2, b2 = c2.csr(256, 256, rp, col, val), c2.empty(4096)
    c2.close()
    b2.free()
    A2.free()
    A = ctx.csr(256, 256, rp, col, val)  # the next call that checks cudaGetLastError
    x, out = ctx.array(np.ones(256, np.float32)), ctx.empty(1024)
    ctx.spmv(A, shb.PLUS_TIMES, x, None, 1.0, 0.0, out)
    want = orc.spmv(orc.PLUS_TIMES, rp, col, val, np.ones(256, np.float32), np.zeros(256, np.float32), 1.0, 0.0)
    assert np.array_equal(out.download(np.float32, 256), want)


@pytest.mark.parametrize("parts,source,window", [(1, 0, 0.0), (3, 0, 0.0), (4, 777, 2.0), (8, 1199, 0.5), (2, 5, 16.0)])
def test_emulated_row_partition_sssp_grid_matches_dijkstra(ctx, orc, parts, source, window):
    """Block-asynchronous near-far on P row blocks of one device (own dist / prev / work per block, candidates for
    other blocks sent as red.min into their dist): bit-exact against the CPU Dijkstra with the same float adds."""
    from sparseharness_b200 import multi
    rp, col, val = orc.gen_grid2d(40, 30)
    n = len(rp) - 1
    A = ctx.csr(n, n, rp, col, val)  # symmetric: rows of A are also the out-edges
    bounds = A.balance_rows(parts)
    blocks = [A.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(parts)]
    be = multi.EmulatedSsspPartition(ctx, blocks, bounds, source, window_buckets=window)
    rounds, conv = multi.run_sssp(be)
    want = orc.sssp(rp, col, val, source)
    assert conv and np.array_equal(be.result(), want)
    assert rounds >= 1 and (parts > 1 or window > 0 or rounds == 1)
    reached, edges = be.stats_total()
    assert reached == n and edges == len(col)
    # the same answer as the single-device driver, and again after a reset
    d1, st = ctx.sssp(A, source, AT=A)
    assert np.array_equal(d1.download(np.float32, n), want)
    be.reset()
    assert multi.run_sssp(be)[1] and np.array_equal(be.result(), want)


def test_emulated_row_partition_sssp_directed_with_unreachable(ctx, orc):
    """Directed weighted operator (out-edges = rows of the transpose), hub rows, unreachable vertices stay FLT_MAX."""
    from sparseharness_b200 import multi
    rng = np.random.default_rng(11)
    rp, col = orc.gen_rmat(11, 8, symmetrise=False)
    n = len(rp) - 1
    val = rng.integers(1, 50, len(col)).astype(np.float32)
    A = ctx.csr(n, n, rp, col, val)
    AT = A.transpose()
    source = int(np.argmax(np.diff(rp)))
    bounds = AT.balance_rows(5)
    at_blocks = [AT.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(5)]
    be = multi.EmulatedSsspPartition(ctx, at_blocks, bounds, source, window_buckets=3.0)
    rounds, conv = multi.run_sssp(be)
    want = orc.sssp(rp, col, val, source)
    assert conv and np.array_equal(be.result(), want)
    assert (want == orc.FLT_MAX).any() and (want < orc.FLT_MAX).sum() > 100


@pytest.mark.parametrize("source,window", [(31000, None), (0, 1000.0), (0, 2.0)])
def test_two_gpu_sssp_matches_oracle(source, window):
    """window 1000 from a corner: rank 1 has nothing to do in round 0 and must still wait for rank 0's candidates
    before its next seed pass (the barrier after the solve)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 30200 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tools", "multi_sssp.py"), "--width", "300",
           "--height", "200", "--source", str(source), "--check"] + (["--window", str(window)] if window is not None else [])
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])


def test_emulated_partitions_with_empty_blocks(ctx, orc):
    """More blocks than 32-vertex words: the trailing blocks own nothing and must still take part in every level /
    round (BFS and SSSP), as a rank with no rows does on a small graph."""
    from sparseharness_b200 import multi
    rp, col, val = orc.gen_grid2d(10, 10)
    n = len(rp) - 1
    A = ctx.csr(n, n, rp, col, val)
    bounds = multi.word_aligned_bounds(np.array([0, 20, 40, 60, 80, 100, 100]))
    assert list(bounds) == [0, 32, 64, 64, 96, 100, 100]
    blocks = [A.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(6)]
    deg = A.row_lengths()
    P = ctx.csr(n, n, rp, col, None)
    pblocks = [P.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(6)]
    be = multi.EmulatedBfsPartition(ctx, pblocks, bounds, 55, len(col), at_blocks=pblocks, deg=deg)
    n_levels, _, _ = multi.run_bfs(be, 1 << 20, 5)
    want, nlev, edges = orc.bfs_levels(rp, col, None, 55)
    assert np.array_equal(be.result(), want) and n_levels == nlev and be.stats() == (n, edges)
    bs = multi.EmulatedSsspPartition(ctx, blocks, bounds, 55, window_buckets=1.0)
    rounds, conv = multi.run_sssp(bs, 4096, 3)
    assert conv and np.array_equal(bs.result(), orc.sssp(rp, col, val, 55))


def test_partition_entry_points_validate_their_arguments(ctx, orc):
    """Bad geometry is SHB_ERR_INVALID (status 1) with a message, never a launch."""
    import sparseharness_b200 as shb
    from sparseharness_b200 import multi
    rp, col, val = orc.gen_grid2d(16, 16)
    n = len(rp) - 1
    A = ctx.csr(n, n, rp, col, val)
    B = A.slice_rows(64, 128)
    d, prev = ctx.empty(n * 4), ctx.empty(128 * 4)
    work = ctx.empty(ctx.sssp_block_work_bytes(128))
    seeds, smin = ctx.empty(64), ctx.empty(64)
    ctx.sssp_block_init(B, 64, n, 0, d, prev, work)
    bad = [
        lambda: ctx.sssp_block_solve(B, 64, n, d, [d], 0, np.array([0, n]), prev, work),             # bounds of another block
        lambda: ctx.sssp_block_solve(B, 64, n, d, [d, d], 1, np.array([0, 64, 200]), prev, work),    # do not end at n
        lambda: ctx.sssp_block_solve(B, 64, n, d, [d], 0, np.array([64, 192]), prev, work),          # do not start at 0
        lambda: ctx.sssp_block_solve(B, 64, n, d, [d, d, d], 1, np.array([0, 64, 192, n]), prev, work, None, 0, 4.0),  # window without seed_min
        lambda: ctx.sssp_block_seed(B, 64, n, d, prev, ctx.empty(16), seeds, smin, 0),               # work too short
        lambda: ctx.sssp_block_seed(B, 64, n, d, prev, work, seeds, smin, 8),                        # no such seed word
        lambda: ctx.sssp_block_init(B, 64, n, n, d, prev, work),                                     # source out of range
        lambda: ctx.sssp_block_init(B, 200, n, 0, d, prev, work),                                    # block outside the vertex set
    ]
    for i, f in enumerate(bad):
        with pytest.raises(shb.ShbError) as e:
            f()
        assert e.value.status == 1, i
    # the good call still works afterwards
    ctx.sssp_block_seed(B, 64, n, d, prev, work, seeds, smin, 0)
    ctx.sssp_block_solve(B, 64, n, d, [d, d, d], 1, np.array([0, 64, 192, n]), prev, work)
    ctx.synchronize()
    # BFS: counters without a slot for the level, a next bitmap aliasing the frontier, wrong self index
    P = ctx.csr(n, n, rp, col, None).slice_rows(64, 128)
    nb = n // 8
    lv, b0, b1, b2, vis, st = ctx.empty(128 * 4), ctx.empty(nb), ctx.empty(nb), ctx.empty(nb), ctx.empty(nb), ctx.empty(ctx.bfs_block_state_bytes())
    cnt = ctx.empty(ctx.bfs_block_counter_bytes()[0])
    cnt.zero()
    ctx.bfs_block_init(P, P, 64, n, len(col), 70, None, lv, b0, b1, b2, vis, cnt, st)
    bad = [
        lambda: ctx.bfs_block_level(P, P, 64, n, 0, None, b0, [b1], 0, b2, vis, lv, ctx.empty(64), st),   # counters too short
        lambda: ctx.bfs_block_level(P, P, 64, n, 0, None, b0, [b0], 0, b2, vis, lv, cnt, st),   # next aliases the frontier
        lambda: ctx.bfs_block_level(P, P, 64, n, 0, None, b0, [b1], 0, b0, vis, lv, cnt, st),   # spare aliases the frontier
        lambda: ctx.bfs_block_level(P, P, 64, n, 0, None, b0, [b1], 1, b2, vis, lv, cnt, st),   # self index out of range
        lambda: ctx.bfs_block_init(P, P, 64, n, len(col), n, None, lv, b0, b1, b2, vis, cnt, st),  # root out of range
    ]
    for i, f in enumerate(bad):
        with pytest.raises(shb.ShbError) as e:
            f()
        assert e.value.status == 1, i
    assert multi.word_aligned_bounds(np.array([0, 64, 192, n]))[-1] == n


def test_emulated_row_partition_bfs_more_levels_than_counter_slots(ctx, orc):
    """A path-like graph with ~1200 levels: the per-level count slots are a ring of 256 frontiers that the traversal
    laps several times (push and pull levels, 3 blocks)."""
    from sparseharness_b200 import multi
    rp, col, _ = orc.gen_grid2d(600, 2)  # 600 x 2 ladder: diameter ~ 600 from a corner
    n = len(rp) - 1
    A = ctx.csr(n, n, rp, col, None)
    bounds = multi.word_aligned_bounds(A.balance_rows(3))
    blocks = [A.slice_rows(int(bounds[k]), int(bounds[k + 1] - bounds[k])) for k in range(3)]
    want, nlev, edges = orc.bfs_levels(rp, col, None, 0)
    assert nlev > 2 * ctx.bfs_block_counter_bytes()[1]
    for push, ce in ((True, 100), (False, 7)):
        be = multi.EmulatedBfsPartition(ctx, blocks, bounds, 0, len(col), at_blocks=blocks if push else None,
                                        deg=A.row_lengths() if push else None)
        n_levels, _, launches = multi.run_bfs(be, 1 << 20, ce)
        assert n_levels == nlev and launches == nlev and np.array_equal(be.result(), want) and be.stats() == (n, edges)
