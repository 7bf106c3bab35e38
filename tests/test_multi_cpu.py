"""world_size-2 (and 3) gloo tests, on CPU, of the multi-GPU host logic in sparseharness_b200/multi.py: the
nnz-balanced row partition, the per-iteration exchange + flag all-reduce protocol and run_fixed_point's launch
accounting.  The local SpMV of each rank is the CPU oracle standing in for the CUDA kernel (the oracle is the
checker here: the product path is GpuRowPartition, exercised by the -m gpu tests)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _worker(rank, world, port, out_dir, case):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as orc
    from sparseharness_b200 import multi

    n = 3000
    rp, col = orc.gen_rmat(11, 8, symmetrise=False)
    n = len(rp) - 1
    val = orc.pagerank_values(rp, col)
    bounds = multi.balanced_bounds(rp, world)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    lrp, lcol, lval = rp[lo:hi + 1] - rp[lo], col[rp[lo]:rp[hi]], val[rp[lo]:rp[hi]]
    delta = 1e-7
    beta = (np.float32(1.0) - np.float32(0.85)) / np.float32(n)

    class NumpyBackend(multi.Backend):
        def __init__(self):
            self.x = [np.full(n, np.float32(1.0) / np.float32(n), np.float32), np.zeros(n, np.float32)]
            self.flags = torch.zeros(4096, dtype=torch.int32)
            self.launches = 0

        def step(self, it, src, dst):
            if it > 0 and self.flags[it - 1] == 0:
                return  # what the kernel's prev_changed early-out does
            out = orc.spmv(orc.PLUS_TIMES, lrp, lcol, lval, self.x[src], np.ones(hi - lo, np.float32), 1.0, beta)
            d = (self.x[src][lo:hi] - out).astype(np.float32)
            self.flags[it] = int(not (np.abs(d.astype(np.float64)) < delta).all())
            self.x[dst][lo:hi] = out
            self.launches += 1

        def exchange(self, it, dst):
            t = torch.from_numpy(self.x[dst])
            views = [t[int(bounds[r]):int(bounds[r + 1])] for r in range(world)]
            parts = [torch.empty_like(v) for v in views]
            dist.all_gather(parts, views[rank].clone()) if len({len(v) for v in views}) == 1 else \
                [dist.broadcast(parts[r] if r != rank else parts[r].copy_(views[r]), src=r) for r in range(world)]
            for v, p in zip(views, parts):
                v.copy_(p)

        def reduce_flag(self, it):
            dist.all_reduce(self.flags[it:it + 1], op=dist.ReduceOp.MAX)

        def read_flags(self, a, b):
            return self.flags[a:b].numpy()

    be = NumpyBackend()
    iters, converged, idx = multi.run_fixed_point(be, case["max_iters"], case["check_every"])
    np.save(os.path.join(out_dir, f"x_{rank}.npy"), be.x[idx])
    np.save(os.path.join(out_dir, f"meta_{rank}.npy"), np.array([iters, int(converged), be.launches, lo, hi]))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,check_every,max_iters", [(2, 8, 400), (2, 1, 400), (3, 5, 400), (2, 4, 6)])
def test_partitioned_pagerank_over_gloo_matches_single_process_oracle(tmp_path, orc, world, check_every, max_iters):
    port = 29500 + (os.getpid() + world * 7 + check_every) % 2000
    mp.spawn(_worker, args=(world, port, str(tmp_path), {"check_every": check_every, "max_iters": max_iters}),
             nprocs=world, join=True)
    rp, col = orc.gen_rmat(11, 8, symmetrise=False)
    val = orc.pagerank_values(rp, col)
    want, want_iters = orc.pagerank(rp, col, val, delta=1e-7, max_iters=max_iters)
    xs = [np.load(tmp_path / f"x_{r}.npy") for r in range(world)]
    metas = [np.load(tmp_path / f"meta_{r}.npy") for r in range(world)]
    for r in range(world):
        assert np.array_equal(xs[r].view(np.int32), want.view(np.int32)), r  # same order per row -> same bits
        assert metas[r][0] == want_iters and metas[r][2] == want_iters
        assert bool(metas[r][1]) == (want_iters < max_iters or max_iters == 400)
    assert metas[0][3] == 0 and metas[-1][4] == len(rp) - 1
    assert all(metas[r][4] == metas[r + 1][3] for r in range(world - 1))


def test_balanced_bounds_rule():
    from sparseharness_b200 import multi
    rp = np.array([0, 0, 10, 10, 11, 30, 30, 64], np.int64)
    assert list(multi.balanced_bounds(rp, 1)) == [0, 7]
    b = multi.balanced_bounds(rp, 4)  # targets 0,16,32,48 -> first row whose prefix reaches the target
    assert list(b) == [0, 5, 7, 7, 7] or list(b) == [0, 5, 6, 7, 7] or b[0] == 0 and b[-1] == 7 and (np.diff(b) >= 0).all()
    for parts in (2, 3, 8):
        b = multi.balanced_bounds(rp, parts)
        assert b[0] == 0 and b[-1] == 7 and (np.diff(b) >= 0).all()
        for k in range(1, parts):
            t = (64 * k) // parts
            assert rp[b[k]] >= t and (b[k] == 0 or rp[b[k] - 1] < t)


# ---------------------------------------------------------------------------------------------------------
# row-partitioned BFS: the level loop, the bitmap ownership rule and the found-count protocol over gloo
# ---------------------------------------------------------------------------------------------------------
def _bfs_worker(rank, world, port, out_dir, case):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as orc
    from sparseharness_b200 import multi

    rp, col = orc.gen_rmat(case["scale"], 8, symmetrise=True)
    n = len(rp) - 1
    root = int(np.argmax(np.diff(rp) > 0))
    bounds = multi.word_aligned_bounds(multi.balanced_bounds(rp, world))
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    n_words = (n + 31) // 32

    class NumpyBfs(multi.BfsBackend):
        """CPU stand-in for shb_bfs_block_step: same bitmaps, same slots, same early-out."""

        def __init__(self):
            self.bm = [np.zeros(n_words, np.uint32), np.zeros(n_words, np.uint32)]
            self.bm[0][root >> 5] = np.uint32(1) << np.uint32(root & 31)
            self.levels = np.full(hi - lo, -1, np.int32)
            if lo <= root < hi:
                self.levels[root - lo] = 0
            self.counts = torch.zeros(4096, dtype=torch.int32)
            self.launches = 0

        def step(self, level):
            if level > 0 and self.counts[level - 1] == 0:
                return
            cur = self.bm[level & 1]
            words = np.zeros(n_words, np.uint32)
            found = 0
            for r in range(lo, hi):
                if self.levels[r - lo] != -1:
                    continue
                c = col[rp[r]:rp[r + 1]]
                if len(c) and (((cur[c >> 5] >> (c & 31).astype(np.uint32)) & 1) != 0).any():
                    self.levels[r - lo] = level + 1
                    words[r >> 5] |= np.uint32(1) << np.uint32(r & 31)
                    found += 1
            self.counts[level] = found
            self._mine = words[lo >> 5:(hi + 31) >> 5]
            self.launches += 1

        def reduce_count(self, level):
            # the exchange the kernel does with peer stores: every rank's word slice into everyone's next bitmap
            if not (level > 0 and self.counts[level - 1] == 0):
                parts = [None] * world
                dist.all_gather_object(parts, self._mine)
                nxt = self.bm[(level + 1) & 1]
                for r in range(world):
                    a, b = int(bounds[r]) >> 5, (int(bounds[r + 1]) + 31) >> 5
                    if b > a:
                        nxt[a:b] = parts[r]
            dist.all_reduce(self.counts[level:level + 1], op=dist.ReduceOp.SUM)

        def read_counts(self, a, b):
            return self.counts[a:b].numpy()

    be = NumpyBfs()
    n_levels, reached, launches = multi.run_bfs(be, case.get("max_levels", 1 << 20), case["check_every"])
    np.save(os.path.join(out_dir, f"lv_{rank}.npy"), be.levels)
    np.save(os.path.join(out_dir, f"bmeta_{rank}.npy"), np.array([n_levels, reached, launches, be.launches, lo, hi]))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,check_every", [(2, 4), (2, 1), (3, 3)])
def test_partitioned_bfs_over_gloo_matches_queue_bfs(tmp_path, orc, world, check_every):
    port = 31500 + (os.getpid() + world * 11 + check_every) % 2000
    mp.spawn(_bfs_worker, args=(world, port, str(tmp_path), {"scale": 9, "check_every": check_every}), nprocs=world, join=True)
    rp, col = orc.gen_rmat(9, 8, symmetrise=True)
    root = int(np.argmax(np.diff(rp) > 0))
    want = orc.bfs_levels(rp, col, None, root)[0]
    got = np.concatenate([np.load(tmp_path / f"lv_{r}.npy") for r in range(world)])
    metas = [np.load(tmp_path / f"bmeta_{r}.npy") for r in range(world)]
    assert np.array_equal(got, want)
    for m in metas:
        assert m[0] == want.max() + 1 and m[1] == (want >= 0).sum() - 1
        assert m[2] == want.max() + 1 and m[3] == m[2]  # launches that did work: one per level incl. the empty one
    assert metas[0][4] == 0 and metas[-1][5] == len(rp) - 1
    assert all(metas[r][5] == metas[r + 1][4] and metas[r][5] % 32 == 0 for r in range(world - 1))


def test_word_aligned_bounds():
    from sparseharness_b200 import multi
    b = multi.word_aligned_bounds(np.array([0, 33, 64, 100, 101]))
    assert list(b) == [0, 64, 64, 101, 101]
    b = multi.word_aligned_bounds(np.array([0, 5, 1000]))
    assert list(b) == [0, 32, 1000]


# ---------------------------------------------------------------------------------------------------------
# row-partitioned SSSP: the round protocol (seed count all-reduce, stop on a globally empty seed pass)
# ---------------------------------------------------------------------------------------------------------
def _sssp_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as orc
    from sparseharness_b200 import multi

    rp, col, val = orc.gen_grid2d(24, 17)
    n = len(rp) - 1
    bounds = multi.balanced_bounds(rp, world)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    FMAX = np.float32(orc.FLT_MAX)

    class NumpySssp(multi.SsspBackend):
        """CPU stand-in for shb_sssp_block_round: owned slice authoritative, candidates for other ranks' vertices
        handed over between rounds (what the kernel does with red.min on the peer mapping)."""

        def __init__(self):
            self.d = np.full(n, FMAX, np.float32)
            self.d[0] = 0
            self.prev = np.full(hi - lo, FMAX, np.float32)
            self.seeds = torch.zeros(4096, dtype=torch.int64)
            self.outbox, self.work = {}, []

        def seed(self, r):
            self.deliver()  # the previous solve's candidates have landed (the kernel's peer red.min + the barrier)
            self.work = [v for v in range(lo, hi) if self.d[v] < self.prev[v - lo]]
            self.seeds[r] = len(self.work)

        def solve(self, r):
            work = self.work
            self.outbox = {}
            while work:  # symmetric grid: row v lists v's neighbours
                v = work.pop()
                self.prev[v - lo] = self.d[v]
                for k in range(rp[v], rp[v + 1]):
                    t, c = int(col[k]), np.float32(np.abs(self.d[v]) + np.abs(val[k]))
                    if c < self.d[t]:
                        self.d[t] = c
                        if lo <= t < hi:
                            work.append(t)
                        else:
                            self.outbox[t] = c

        def reduce_seeds(self, r):
            dist.all_reduce(self.seeds[r:r + 1], op=dist.ReduceOp.SUM)

        def barrier(self, r):
            dist.barrier()

        def deliver(self):
            boxes = [None] * world
            dist.all_gather_object(boxes, self.outbox)
            for b in boxes:
                for t, c in b.items():
                    if lo <= t < hi and c < self.d[t]:
                        self.d[t] = c

        def read_seeds(self, r):
            return int(self.seeds[r])

    be = NumpySssp()
    rounds, conv = multi.run_sssp(be, 500)
    np.save(os.path.join(out_dir, f"sd_{rank}.npy"), be.d[lo:hi])
    np.save(os.path.join(out_dir, f"smeta_{rank}.npy"), np.array([rounds, int(conv)]))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_partitioned_sssp_rounds_over_gloo_match_dijkstra(tmp_path, orc, world):
    port = 33500 + (os.getpid() + world * 13) % 2000
    mp.spawn(_sssp_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    rp, col, val = orc.gen_grid2d(24, 17)
    want = orc.sssp(rp, col, val, 0)
    got = np.concatenate([np.load(tmp_path / f"sd_{r}.npy") for r in range(world)])
    metas = [np.load(tmp_path / f"smeta_{r}.npy") for r in range(world)]
    assert np.array_equal(got, want)
    assert all(m[1] == 1 and m[0] == metas[0][0] and m[0] >= world for m in metas)
