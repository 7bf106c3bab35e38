"""The bounded-cost parity gates bench.py runs at full size (tools/parity_checks.py) accept the oracle's answers
and reject perturbed ones."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import parity_checks as pc  # noqa: E402


def test_bfs_local_conditions_characterise_the_levels(orc):
    rp, col = orc.gen_rmat(11, 8, seed=orc.SEED_C3, symmetrise=True)
    n = len(rp) - 1
    root = int(np.argmax(np.diff(rp) > 0))
    lv, _, _ = orc.bfs_levels(rp, col, None, root)
    assert (lv < 0).any() and (lv > 1).any()
    begins, cnt = pc.sample_blocks(n, 256, extra=[700])
    for b in begins:
        sub = rp[b:b + cnt + 1] - rp[b]
        c = col[rp[b]:rp[b + cnt]]
        assert pc.bfs_rows_ok(lv, sub, c, b, root)
    assert pc.bfs_rows_ok(lv, rp, col, 0, root)
    for v, delta in ((int(np.argmax(lv == 2)), 1), (int(np.argmax(lv == 2)), -1), (int(np.argmax(lv == 1)), -2)):
        bad = lv.copy()
        bad[v] += delta
        assert not pc.bfs_rows_ok(bad, rp, col, 0, root)
    bad = lv.copy()
    bad[int(np.argmax((lv < 0) & (np.diff(rp) > 0)))] = 3  # a vertex of another component marked reached
    assert not pc.bfs_rows_ok(bad, rp, col, 0, root)


def test_bellman_condition_characterises_the_distances(orc):
    rp, col, val = orc.gen_grid2d(48, 40)
    d = orc.sssp(rp, col, val, 0)
    assert pc.sssp_rows_ok(d, rp, col, val, 0, 0)
    b = 700
    assert pc.sssp_rows_ok(d, rp[b:b + 101] - rp[b], col[rp[b]:rp[b + 100]], val[rp[b]:rp[b + 100]], b, 0)
    for delta in (1.0, -1.0):
        bad = d.copy()
        bad[777] += np.float32(delta)
        assert not pc.sssp_rows_ok(bad, rp, col, val, 0, 0)
    assert pc.checksum64(d) != pc.checksum64(bad) and pc.checksum64(d) == pc.checksum64(d[::-1].copy())


def test_pagerank_sweep_check(orc):
    rp, col = orc.gen_rmat(10, 8, seed=orc.SEED_C3, symmetrise=False)
    val = orc.pagerank_values(rp, col)
    n = len(rp) - 1
    beta = float((np.float32(1.0) - np.float32(0.85)) / np.float32(n))
    x19, _ = orc.pagerank(rp, col, val, delta=0.0, max_iters=19)
    x20, _ = orc.pagerank(rp, col, val, delta=0.0, max_iters=20)
    assert pc.pagerank_rows_err(orc, x19, x20, rp, col, val, beta) <= 1e-6
    bad = x20.copy()
    bad[5] *= np.float32(1.01)
    assert pc.pagerank_rows_err(orc, x19, bad, rp, col, val, beta) > 1e-5
