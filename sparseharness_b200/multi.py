"""Multi-GPU fixed-point iteration of the semiring SpMV: 1-D row partition, x exchanged every iteration.

New capability (the reference is single-device, inc/harness.h:78-81); specified by BASELINE.json's
north_star and SURVEY.md 8e.  One process per GPU (torchrun), torch.distributed for rendezvous and
collectives.  Rank r owns the contiguous row block [bounds[r], bounds[r+1]) of the operator -- blocks are
balanced by non-zeros -- with GLOBAL column ids, and two full-length x buffers (ping-pong).

Per iteration, on every rank:
    1. local SpMV of the row block from x[cur] (skipped on the device when the previous iteration's
       all-reduced flag says the fixed point was reached),
    2. exchange so that every rank's x[next] holds all blocks:
         "nccl"  : the kernel writes its own block into the local x[next]; the blocks are all-gathered over NCCL
         "fused" : shb_spmv_scatter's epilogue stores every finished out[r] into ALL ranks' x[next] through
                   CUDA-IPC mapped peer pointers over NVLink -- compute and exchange are one kernel,
    3. a 4-byte all-reduce(MAX) of the block's "some row still changes" flag: convergence test and, being
       stream ordered after the kernel on every rank, the barrier that orders peer writes before the next read.
The host looks at the flags once per `check_every` iterations; nothing else leaves the device.

The loop (`run_fixed_point`) is backend agnostic: tests drive it on CPU with a numpy backend over gloo.
"""
from __future__ import annotations

import numpy as np


def balanced_bounds(row_ptr: np.ndarray, parts: int) -> np.ndarray:
    """bounds[k] = first row r with row_ptr[r] >= k*nnz/parts (the rule of shb_csr_balance_rows)."""
    row_ptr = np.asarray(row_ptr, dtype=np.int64)
    n, nnz = len(row_ptr) - 1, int(row_ptr[-1])
    b = np.empty(parts + 1, dtype=np.int64)
    for k in range(parts + 1):
        b[k] = n if k == parts else (0 if k == 0 else int(np.searchsorted(row_ptr[:n], (nnz * k) // parts, side="left")))
    return b


class Backend:
    """What run_fixed_point needs from a device / a CPU stand-in.  Buffers are indexed 0/1 (ping-pong)."""

    def step(self, it: int, src: int, dst: int) -> None:  # local SpMV (+ exchange when fused) of iteration `it`
        raise NotImplementedError

    def exchange(self, it: int, dst: int) -> None:  # make x[dst] complete on every rank
        raise NotImplementedError

    def reduce_flag(self, it: int) -> None:  # flags[it] <- max over ranks (stream ordered, no host sync)
        raise NotImplementedError

    def read_flags(self, lo: int, hi: int) -> np.ndarray:  # host view of flags[lo:hi] (synchronises)
        raise NotImplementedError

    def begin_window(self, lo: int, hi: int) -> None:  # optional hook before a batch of iterations
        pass


def run_fixed_point(backend: Backend, max_iters: int, check_every: int = 8):
    """x_{k+1} = KERNEL(x_k) until no row changes on any rank.  Returns (iterations, converged, result_index):
    iterations = launches that did work (the converging one included), result_index = which of the two x
    buffers holds the last output.  Mirrors shb_iterate (shb_core.cu) launch for launch."""
    assert max_iters >= 1 and check_every >= 1
    done, cur, converged = 0, 0, False
    while done < max_iters and not converged:
        batch = min(check_every, max_iters - done)
        backend.begin_window(done, done + batch)
        for b in range(batch):
            it = done + b
            backend.step(it, cur, cur ^ 1)
            backend.exchange(it, cur ^ 1)
            backend.reduce_flag(it)
            cur ^= 1
        flags = backend.read_flags(done, done + batch)
        ran = batch
        for b in range(batch):
            if flags[b] == 0:  # iteration done+b changed nothing anywhere: fixed point; later launches returned at once
                ran, converged = b + 1, True
                break
        if (batch - ran) & 1:
            cur ^= 1
        done += ran
    return done, converged, cur


class GpuRowPartition(Backend):
    """The CUDA backend: this rank's row block on its GPU."""

    def __init__(self, ctx, A_local, bounds, rank, world, semiring, alpha, beta, *, mode="fused", y_mode="one",
                 y0=None, delta=1e-4, term=0, cfg=None, max_iters=1024, dist=None, torch=None):
        import sparseharness_b200 as shb
        self.shb, self.ctx, self.A, self.bounds = shb, ctx, A_local, np.asarray(bounds, dtype=np.int64)
        self.rank, self.world, self.semiring, self.alpha, self.beta = rank, world, semiring, alpha, beta
        self.mode, self.y_mode, self.delta, self.term = mode, y_mode, delta, term
        self.dist, self.torch = dist, torch
        self.n = int(self.bounds[-1])
        self.lo, self.hi = int(self.bounds[rank]), int(self.bounds[rank + 1])
        flags = cfg.flags if cfg is not None else 0
        if y_mode == "one":
            flags |= shb.F_Y_IS_ONE
        self.cfg = shb.launch_cfg(cfg.variant if cfg else 0, cfg.lanes if cfg else 0, cfg.rows_per_group if cfg else 0,
                                  cfg.block_threads if cfg else 0, cfg.grid_blocks if cfg else 0, flags)
        self.y0 = y0  # Buffer, full length, used when y_mode == "fixed" (or for iteration 0 of "rebind")
        self.x = [ctx.empty(max(self.n, 1) * 4), ctx.empty(max(self.n, 1) * 4)]
        self.max_flags = max_iters + 1
        self.flags = ctx.empty(self.max_flags * 4)
        self.flags.zero()
        self._slot = [ctx.wrap(self.flags.device_ptr + 4 * i, 4) for i in range(self.max_flags)]
        self.launches = 0
        self.peers = None
        if world > 1:
            self._t_x = [self._as_tensor(b, self.n, torch.float32 if semiring < 2 else torch.int32) for b in self.x]
            self._t_flags = self._as_tensor(self.flags, self.max_flags, torch.int32)
            if mode == "fused":
                self._open_peers()

    # ---- plumbing
    def _as_tensor(self, buf, count, dtype):
        """A torch view of a libshb buffer (for the NCCL collectives only)."""
        torch = self.torch

        class _Cai:
            pass

        o = _Cai()
        o.__cuda_array_interface__ = {"shape": (count,), "typestr": "<f4" if dtype == torch.float32 else "<i4",
                                      "data": (buf.device_ptr, False), "version": 2}
        return torch.as_tensor(o, device=f"cuda:{self.ctx.device}")

    def _open_peers(self):
        handles = [None] * self.world
        self.dist.all_gather_object(handles, (self.x[0].ipc_handle(), self.x[1].ipc_handle()))
        self.peers = [[], []]
        for r in range(self.world):
            for k in (0, 1):
                self.peers[k].append(self.x[k] if r == self.rank else self.ctx.ipc_open(handles[r][k], self.n * 4))

    def set_x0(self, fill_bits=None, host=None):
        if host is not None:
            self.x[0].upload(host)
        else:
            self.x[0].fill_u32(fill_bits, self.n)

    # ---- Backend
    def step(self, it, src, dst):
        y = None
        if self.y_mode == "fixed" or (self.y_mode == "rebind" and it == 0):
            y = self.y0
        dsts = self.peers[dst] if (self.mode == "fused" and self.world > 1) else [self.x[dst]]
        self.ctx.spmv_scatter(self.A, self.semiring, self.x[src], y, self.alpha, self.beta, dsts, self.lo,
                              delta=self.delta, term=self.term, changed=self._slot[it],
                              prev_changed=self._slot[it - 1] if it > 0 else None, cfg=self.cfg)
        self.launches += 1

    def exchange(self, it, dst):
        if self.world == 1 or self.mode == "fused":
            return
        t = self._t_x[dst]
        sizes = np.diff(self.bounds)
        if (sizes == sizes[0]).all():  # equal blocks: one in-place ncclAllGather
            self.dist.all_gather_into_tensor(t, t[self.lo:self.hi])
        else:  # uneven blocks: torch issues one grouped ncclBroadcast per block straight into the views
            views = [t[int(self.bounds[r]):int(self.bounds[r + 1])] for r in range(self.world)]
            self.dist.all_gather(views, views[self.rank])

    def reduce_flag(self, it):
        if self.world > 1:
            self.dist.all_reduce(self._t_flags[it:it + 1], op=self.dist.ReduceOp.MAX)

    def read_flags(self, lo, hi):
        return self.flags.download(np.uint32, hi - lo, offset=4 * lo)

    def result(self, index, dtype=np.float32):
        return self.x[index].download(dtype, self.n)

    def close(self):
        if self.peers:
            for k in (0, 1):
                for r, b in enumerate(self.peers[k]):
                    if r != self.rank:
                        b.free()
        for b in self.x + [self.flags]:
            b.free()


def partitioned_pagerank(ctx, A_local, bounds, rank, world, *, damping=0.85, delta=1e-4, max_iters=1000, check_every=8,
                         mode="fused", cfg=None, dist=None, torch=None):
    """Canonical PageRank on a row-partitioned operator whose values already hold d / colsum.
    Returns (backend, iterations, converged, result_index, device_ms)."""
    import sparseharness_b200 as shb
    n = int(bounds[-1])
    x_init = np.float32(1.0) / np.float32(n)
    beta = (np.float32(1.0) - np.float32(damping)) / np.float32(n)
    be = GpuRowPartition(ctx, A_local, bounds, rank, world, shb.PLUS_TIMES, 1.0, float(beta), mode=mode, y_mode="one",
                         delta=delta, term=shb.TERM_ABSDIFF, cfg=cfg, max_iters=max_iters, dist=dist, torch=torch)
    be.set_x0(fill_bits=int(np.array([x_init], np.float32).view(np.uint32)[0]))
    ctx.synchronize()
    if world > 1:
        dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    iters, converged, idx = run_fixed_point(be, max_iters, check_every)
    ev1.record()
    torch.cuda.synchronize()
    return be, iters, converged, idx, ev0.elapsed_time(ev1)


class EmulatedRowPartition(Backend):
    """All P row blocks on ONE device: one scatter launch per block and iteration, every block writing its
    slice of the shared x[next].  The arithmetic, offsets and flag protocol are exactly the multi-GPU path's;
    used where fewer GPUs than ranks are available (tests, B200_PROFILING.md's advice for 1-GPU boxes)."""

    def __init__(self, ctx, blocks, bounds, semiring, alpha, beta, *, y_mode="one", y0=None, delta=1e-4, term=0,
                 cfg=None, max_iters=1024):
        import sparseharness_b200 as shb
        self.ctx, self.blocks, self.bounds = ctx, blocks, np.asarray(bounds, dtype=np.int64)
        self.semiring, self.alpha, self.beta, self.y_mode, self.y0 = semiring, alpha, beta, y_mode, y0
        self.delta, self.term = delta, term
        self.n = int(self.bounds[-1])
        flags = (cfg.flags if cfg is not None else 0) | (shb.F_Y_IS_ONE if y_mode == "one" else 0)
        self.cfg = shb.launch_cfg(cfg.variant if cfg else 0, cfg.lanes if cfg else 0, cfg.rows_per_group if cfg else 0,
                                  cfg.block_threads if cfg else 0, cfg.grid_blocks if cfg else 0, flags)
        self.x = [ctx.empty(max(self.n, 1) * 4), ctx.empty(max(self.n, 1) * 4)]
        self.flags = ctx.empty((max_iters + 1) * 4)
        self.flags.zero()
        self._slot = [ctx.wrap(self.flags.device_ptr + 4 * i, 4) for i in range(max_iters + 1)]
        self.launches = 0

    def step(self, it, src, dst):
        y = self.y0 if (self.y_mode == "fixed" or (self.y_mode == "rebind" and it == 0)) else None
        for k, A in enumerate(self.blocks):
            self.ctx.spmv_scatter(A, self.semiring, self.x[src], y, self.alpha, self.beta, [self.x[dst]],
                                  int(self.bounds[k]), delta=self.delta, term=self.term, changed=self._slot[it],
                                  prev_changed=self._slot[it - 1] if it > 0 else None, cfg=self.cfg)
            self.launches += 1

    def exchange(self, it, dst):
        pass

    def reduce_flag(self, it):
        pass

    def read_flags(self, lo, hi):
        return self.flags.download(np.uint32, hi - lo, offset=4 * lo)

    def result(self, index, dtype=np.float32):
        return self.x[index].download(dtype, self.n)


# =====================================================================================================
# Row-partitioned BFS (integer levels under or-and)
# =====================================================================================================
def word_aligned_bounds(bounds: np.ndarray, align: int = 32) -> np.ndarray:
    """Row bounds rounded to multiples of `align` (the last one stays n): a 32-vertex word of the frontier
    bitmap then has exactly one owner, which is what shb_bfs_block_step requires."""
    b = np.asarray(bounds, dtype=np.int64).copy()
    n = int(b[-1])
    for k in range(1, len(b) - 1):
        b[k] = min(n, (int(b[k]) + align - 1) // align * align)
    for k in range(1, len(b)):
        b[k] = max(b[k], b[k - 1])
    return b


class BfsBackend:
    """What run_bfs needs from a device / CPU stand-in."""

    def step(self, level: int) -> None:  # pull launch of `level` (+ the stores into every rank's next bitmap)
        raise NotImplementedError

    def reduce_count(self, level: int) -> None:  # found[level] <- sum over ranks (stream ordered; also the barrier)
        raise NotImplementedError

    def read_counts(self, lo: int, hi: int) -> np.ndarray:  # host view of found[lo:hi] (synchronises)
        raise NotImplementedError

    def label_frontier(self, level: int) -> None:  # level cap reached: the vertices of frontier `level` get their level
        pass


def run_bfs(backend: BfsBackend, max_levels: int = 1 << 20, check_every: int = 4):
    """Level-synchronous loop: level L's launch builds frontier L+1; stops after the first level that finds
    nothing.  Returns (levels_populated, reached_beyond_root, launches_that_did_work).  Launches issued past
    the end of the traversal inside a batch return at once on the device (their found slot stays 0)."""
    level, reached, done = 0, 0, False
    check_every = max(1, min(check_every, 128))  # the count slots are a ring (256 frontiers): poll well inside it
    while not done and level < max_levels:
        batch = min(check_every, max_levels - level)
        for b in range(batch):
            backend.step(level + b)
            backend.reduce_count(level + b)
        counts = backend.read_counts(level, level + batch)
        for b in range(batch):
            if counts[b] == 0:
                level += b + 1
                done = True
                break
            reached += int(counts[b])
        else:
            level += batch
    if not done:  # cut by the cap: a push level leaves the labelling of the frontier it built to the next level's prep
        backend.label_frontier(level)
    # `level` launches did work; the last of them found nothing when done, so levels 0..level-1 are populated
    return (level if done else level + 1), reached, level


_BFS_SLOT0 = 6  # int64 word of counter slot 0 (include/shb.h: slot F = words 6 + 2 (F mod ring), + 1)


class GpuBfsPartition(BfsBackend):
    """The CUDA backend: this rank's row block.  AT_local (rows = out-edges of the owned vertices) enables push
    levels; deg (full-length int32 push degrees) steers the direction heuristic.  exchange="peer": the per-level
    counts are added into every rank's counters by a publish kernel over CUDA-IPC peer pointers, which also is the
    barrier -- no collective library inside the traversal; exchange="nccl": a 16-byte all-reduce per level."""

    def __init__(self, ctx, A_local, bounds, rank, world, root, nnz_total, *, AT_local=None, deg=None, max_levels=1 << 30,
                 exchange="peer", dist=None, torch=None):
        self.ctx, self.A, self.AT, self.deg = ctx, A_local, AT_local, deg
        self.bounds = np.asarray(bounds, dtype=np.int64)
        self.rank, self.world, self.root, self.dist, self.torch = rank, world, int(root), dist, torch
        self.exchange = exchange if world > 1 else "none"
        self.n, self.nnz_total = int(self.bounds[-1]), int(nnz_total)
        self.lo, self.hi = int(self.bounds[rank]), int(self.bounds[rank + 1])
        self.n_words = max((self.n + 31) // 32, 1)
        self.max_levels = max_levels
        self.bm = [ctx.empty(self.n_words * 4) for _ in range(3)]  # frontier L lives in bitmap L % 3
        self.visited = ctx.empty(self.n_words * 4)
        self.levels = ctx.empty(max(self.hi - self.lo, 1) * 4)
        nbytes, self.ring = ctx.bfs_block_counter_bytes()
        self.n_counter_words = nbytes // 8
        self.counters = ctx.empty(nbytes)
        self.counters.zero()  # the arrival word starts at 0 and is never reset afterwards
        self.state = ctx.empty(ctx.bfs_block_state_bytes())
        self.launches, self.seq = 0, 0
        self.peers = [[self.bm[0]], [self.bm[1]], [self.bm[2]]]
        self.peer_counters = None
        self.self_index = 0
        self._t_counts = None
        if world > 1:
            ctx.synchronize()
            self._t_counts = _tensor_view(torch, self.counters, self.n_counter_words, ctx.device, "<i8")
            handles = [None] * world
            dist.all_gather_object(handles, tuple(b.ipc_handle() for b in self.bm) + (self.counters.ipc_handle(),))
            self.peers = [[], [], []]
            for r in range(world):
                for k in range(3):
                    self.peers[k].append(self.bm[k] if r == rank else ctx.ipc_open(handles[r][k], self.n_words * 4))
            if self.exchange == "peer":
                self.peer_counters = [self.counters if r == rank else ctx.ipc_open(handles[r][3], self.n_counter_words * 8)
                                      for r in range(world)]
            self.self_index = rank
        self.reset()

    def _quiesce(self):
        self.ctx.synchronize()
        if self.world > 1:
            self.dist.barrier()

    def reset(self):
        self._quiesce()  # no peer may still be storing into what init is about to zero
        peer = self.exchange == "peer"
        self.ctx.bfs_block_init(self.A, self.AT, self.lo, self.n, self.nnz_total, self.root, self.deg, self.levels,
                                self.bm[0], self.bm[1], self.bm[2], self.visited, self.counters, self.state, staged=peer)
        if peer:
            self._quiesce()  # every rank has zeroed its slots before anyone adds to them
            self.seq += 1
            self.ctx.bfs_block_publish(0, self.counters, self.state, self.peer_counters, self.self_index,
                                       self.world * self.seq)
        elif self.exchange == "nccl":
            self.dist.all_reduce(self._t_counts[_BFS_SLOT0:_BFS_SLOT0 + 2], op=self.dist.ReduceOp.SUM)
        self.launches = 0

    def step(self, level):
        target = 0
        if self.exchange == "peer":
            self.seq += 1
            target = self.world * self.seq
        self.ctx.bfs_block_level(self.A, self.AT, self.lo, self.n, level, self.deg, self.bm[level % 3],
                                 self.peers[(level + 1) % 3], self.self_index, self.bm[(level + 2) % 3], self.visited,
                                 self.levels, self.counters, self.state, self.peer_counters, target)
        self.launches += 1

    def label_frontier(self, level):
        self.ctx.bfs_block_label_frontier(self.A, self.lo, self.n, level, self.bm[level % 3], self.bm[(level + 2) % 3], self.visited,
                                          self.levels, self.counters, self.state)

    def reduce_count(self, level):
        if self.exchange == "nccl":
            w = _BFS_SLOT0 + 2 * ((level + 1) % self.ring)
            self.dist.all_reduce(self._t_counts[w:w + 2], op=self.dist.ReduceOp.SUM)

    def read_counts(self, lo, hi):
        w = self.counters.download(np.int64, self.n_counter_words)
        return np.array([w[_BFS_SLOT0 + 2 * ((f + 1) % self.ring)] for f in range(lo, hi)], dtype=np.int64)

    def local_stats(self):
        """(reached, sum of in-degrees of reached) of the owned slice; raises if a peer barrier timed out."""
        self.ctx.bfs_block_finish(self.A, self.levels, self.counters, self.state)
        w = self.counters.download(np.int64, 2)
        return int(w[0]), int(w[1])

    def result(self) -> np.ndarray:
        return self.levels.download(np.int32, self.hi - self.lo)

    def close(self):
        self._quiesce()
        if self.world > 1:
            for k in range(3):
                for r, b in enumerate(self.peers[k]):
                    if r != self.rank:
                        b.free()
            for r, b in enumerate(self.peer_counters or []):
                if r != self.rank:
                    b.free()
        for b in self.bm + [self.visited, self.levels, self.counters, self.state]:
            b.free()


def _tensor_view(torch, buf, count, device, typestr="<i4"):
    """A torch view of a libshb buffer (for the NCCL collectives only)."""

    class _Cai:
        pass

    o = _Cai()
    o.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr, "data": (buf.device_ptr, False), "version": 2}
    return torch.as_tensor(o, device=f"cuda:{device}")


class EmulatedBfsPartition(BfsBackend):
    """All P row blocks on ONE device, each with its OWN bitmaps, state and counters, every block storing into every
    block's next bitmap: the multi-GPU path's kernels and protocol on one GPU.  publish=True runs the peer publish
    kernel (system-scope adds into every block's slot, no wait: the blocks share a stream); publish=False emulates
    the NCCL all-reduce of a slot through the host (download, sum, upload)."""

    def __init__(self, ctx, blocks, bounds, root, nnz_total, *, at_blocks=None, deg=None, max_levels=1 << 30, publish=True):
        self.ctx, self.blocks, self.at_blocks, self.deg, self.publish = ctx, blocks, at_blocks, deg, publish
        self.bounds, self.root, self.nnz_total = np.asarray(bounds, dtype=np.int64), int(root), int(nnz_total)
        self.n = int(self.bounds[-1])
        self.P = len(blocks)
        n_words = max((self.n + 31) // 32, 1)
        self.bm = [[ctx.empty(n_words * 4) for _ in range(3)] for _ in range(self.P)]
        self.visited = [ctx.empty(n_words * 4) for _ in range(self.P)]
        self.levels = [ctx.empty(max(int(self.bounds[k + 1] - self.bounds[k]), 1) * 4) for k in range(self.P)]
        nbytes, self.ring = ctx.bfs_block_counter_bytes()
        self.n_counter_words = nbytes // 8
        self.counters = [ctx.empty(nbytes) for _ in range(self.P)]
        for c in self.counters:
            c.zero()
        self.state = [ctx.empty(ctx.bfs_block_state_bytes()) for _ in range(self.P)]
        self.reset()

    def _reduce_slot(self, slot):
        off = 8 * (_BFS_SLOT0 + 2 * (slot % self.ring))
        tot = sum(c.download(np.int64, 2, offset=off) for c in self.counters)
        for c in self.counters:
            c.upload(tot, offset=off)

    def reset(self):
        for k, A in enumerate(self.blocks):
            self.ctx.bfs_block_init(A, self.at_blocks[k] if self.at_blocks else None, int(self.bounds[k]), self.n,
                                    self.nnz_total, self.root, self.deg, self.levels[k], self.bm[k][0], self.bm[k][1],
                                    self.bm[k][2], self.visited[k], self.counters[k], self.state[k], staged=self.publish)
        if self.publish:
            for k in range(self.P):
                self.ctx.bfs_block_publish(0, self.counters[k], self.state[k], self.counters, k, 0)
        else:
            self._reduce_slot(0)
        self.launches = 0

    def step(self, level):
        nxt = [self.bm[k][(level + 1) % 3] for k in range(self.P)]
        for k, A in enumerate(self.blocks):
            self.ctx.bfs_block_level(A, self.at_blocks[k] if self.at_blocks else None, int(self.bounds[k]), self.n, level,
                                     self.deg, self.bm[k][level % 3], nxt, k, self.bm[k][(level + 2) % 3], self.visited[k],
                                     self.levels[k], self.counters[k], self.state[k],
                                     self.counters if self.publish else None, 0)
            self.launches += 1

    def label_frontier(self, level):
        for k, A in enumerate(self.blocks):
            self.ctx.bfs_block_label_frontier(A, int(self.bounds[k]), self.n, level, self.bm[k][level % 3], self.bm[k][(level + 2) % 3],
                                              self.visited[k], self.levels[k], self.counters[k], self.state[k])

    def reduce_count(self, level):
        if not self.publish:
            self._reduce_slot(level + 1)

    def read_counts(self, lo, hi):
        w = self.counters[0].download(np.int64, self.n_counter_words)
        return np.array([w[_BFS_SLOT0 + 2 * ((f + 1) % self.ring)] for f in range(lo, hi)], dtype=np.int64)

    def stats(self):
        r = e = 0
        for k, A in enumerate(self.blocks):
            self.ctx.bfs_block_finish(A, self.levels[k], self.counters[k], self.state[k])
            w = self.counters[k].download(np.int64, 2)
            r, e = r + int(w[0]), e + int(w[1])
        return r, e

    def result(self) -> np.ndarray:
        return np.concatenate([self.levels[k].download(np.int32, int(self.bounds[k + 1] - self.bounds[k]))
                               for k in range(self.P)])


def partitioned_bfs(ctx, A_local, bounds, rank, world, root, nnz_total, *, AT_local=None, deg=None, check_every=8,
                    max_levels=1 << 30, exchange="peer", dist=None, torch=None, repeat=1):
    """Direction-optimising BFS levels on a row-partitioned operator.  Returns (backend, stats) with stats =
    dict(levels, reached, edges_traversed [sum of in-degrees of reached vertices, all ranks], ms [device time of
    this rank; max over ranks is the caller's job], launches)."""
    be = GpuBfsPartition(ctx, A_local, bounds, rank, world, root, nnz_total, AT_local=AT_local, deg=deg,
                         max_levels=max_levels, exchange=exchange, dist=dist, torch=torch)
    best = None
    for i in range(max(1, repeat)):
        if i:
            be.reset()
        ctx.synchronize()
        if world > 1:
            dist.barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        n_levels, _, launches = run_bfs(be, max_levels, check_every)
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        if best is None or ms < best["ms"]:
            best = {"levels": n_levels, "ms": ms, "launches": launches}
    reached, edges = be.local_stats()
    if world > 1:
        t = torch.tensor([reached, edges], dtype=torch.int64, device=f"cuda:{ctx.device}")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        reached, edges = int(t[0].item()), int(t[1].item())
    best["reached"], best["edges_traversed"] = reached, edges
    return be, best


# =====================================================================================================
# Row-partitioned SSSP (min-plus to the exact fixed point): block-asynchronous near-far
# =====================================================================================================
class SsspBackend:
    def seed(self, r: int) -> None:  # seed pass of round r: pending owned vertices, their count and smallest distance
        raise NotImplementedError

    def reduce_seeds(self, r: int) -> None:  # seeds[r] <- sum, seed_min[r] <- min over ranks; the barrier between rounds
        raise NotImplementedError

    def solve(self, r: int) -> None:  # local near-far solve of round r inside the distance window (peer reds included)
        raise NotImplementedError

    def barrier(self, r: int) -> None:  # every rank's solve of round r (and its peer reds) is complete
        raise NotImplementedError

    def read_seeds(self, r: int) -> int:  # host view of seeds[r] (synchronises)
        raise NotImplementedError


def run_sssp(backend: SsspBackend, max_rounds: int = 1 << 16, check_every: int = 8):
    """Rounds until a seed pass finds nothing on any rank (then that round's solves did nothing either: fixed point).
    `check_every` rounds are enqueued per host poll, so the host's launch work overlaps the device's rounds; rounds
    enqueued past the fixed point find no seeds and return at once.  Returns (rounds_with_work, converged)."""
    done = 0
    while done < max_rounds:
        batch = min(check_every, max_rounds - done)
        for r in range(done, done + batch):
            backend.seed(r)
            backend.reduce_seeds(r)  # global seed count and smallest pending distance of this round
            backend.solve(r)         # a solve without seeds returns at once
            backend.barrier(r)       # all peer candidates of this round have landed before anyone's next seed pass
        for r in range(done, done + batch):
            if backend.read_seeds(r) == 0:
                return r, True
        done += batch
    return max_rounds, False


class GpuSsspPartition(SsspBackend):
    """The CUDA backend: AT_local = out-edges of the owned vertices (global target ids)."""

    def __init__(self, ctx, AT_local, bounds, rank, world, source, *, max_rounds=4096, window_buckets=None, dist=None,
                 torch=None):
        self.ctx, self.AT, self.bounds = ctx, AT_local, np.asarray(bounds, dtype=np.int64)
        # one rank: no window (a single solve).  Several ranks: sweep the distance range in windows of this many
        # near-far buckets so that no block runs far ahead on boundary distances a neighbour will still lower.
        self.window = float((0 if world == 1 else 8) if window_buckets is None else window_buckets)
        self.rank, self.world, self.source, self.dist_mod, self.torch = rank, world, int(source), dist, torch
        self.n = int(self.bounds[-1])
        self.lo, self.hi = int(self.bounds[rank]), int(self.bounds[rank + 1])
        self.d = ctx.empty(max(self.n, 1) * 4)
        self.prev = ctx.empty(max(self.hi - self.lo, 1) * 4)
        self.work = ctx.empty(ctx.sssp_block_work_bytes(self.hi - self.lo))
        self.max_rounds = max_rounds
        self.seeds = ctx.empty((max_rounds + 2) * 8)
        self.smin = ctx.empty((max_rounds + 2) * 8)
        self.stats = ctx.empty(16)
        self.peer_d = [self.d]
        self._t_seeds = self._t_smin = None
        if world > 1:
            ctx.synchronize()
            self._t_seeds = _tensor_view(torch, self.seeds, max_rounds + 2, ctx.device, "<i8")
            self._t_smin = _tensor_view(torch, self.smin, max_rounds + 2, ctx.device, "<i8")
            handles = [None] * world
            dist.all_gather_object(handles, self.d.ipc_handle())
            self.peer_d = [self.d if r == rank else ctx.ipc_open(handles[r], self.n * 4) for r in range(world)]
        self.reset()

    def _quiesce(self):
        self.ctx.synchronize()
        if self.world > 1:
            self.dist_mod.barrier()

    def reset(self):
        self._quiesce()  # no peer may still be sending candidates into the dist that init is about to overwrite
        self.seeds.zero()
        self.smin.fill_u32(0x7F7FFFFF)  # as int64 words: larger than any distance
        self.ctx.sssp_block_init(self.AT, self.lo, self.n, self.source, self.d, self.prev, self.work)
        self._quiesce()

    def seed(self, r):
        self.ctx.sssp_block_seed(self.AT, self.lo, self.n, self.d, self.prev, self.work, self.seeds, self.smin, r)

    def reduce_seeds(self, r):
        if self.world > 1:
            self.dist_mod.all_reduce(self._t_seeds[r:r + 1], op=self.dist_mod.ReduceOp.SUM)
            if self.window > 0:
                self.dist_mod.all_reduce(self._t_smin[r:r + 1], op=self.dist_mod.ReduceOp.MIN)

    def solve(self, r):
        self.ctx.sssp_block_solve(self.AT, self.lo, self.n, self.d, self.peer_d, self.rank if self.world > 1 else 0,
                                  self.bounds, self.prev, self.work, self.smin, r, self.window)

    def barrier(self, r):
        if self.world > 1:  # stream-ordered after the solve on every rank; MAX over identical values changes nothing
            self.dist_mod.all_reduce(self._t_seeds[r:r + 1], op=self.dist_mod.ReduceOp.MAX)

    def read_seeds(self, r):
        return int(self.seeds.download(np.int64, 1, offset=8 * r)[0])

    def local_stats(self, A_rows=None):
        """(reached, sum of in-degrees of reached) of the owned slice; raises if the last solve failed."""
        self.ctx.sssp_block_finish(A_rows or self.AT, self.lo, self.d, self.work, self.stats)
        w = self.stats.download(np.int64, 2)
        return int(w[0]), int(w[1])

    def result(self) -> np.ndarray:
        return self.d.download(np.float32, self.hi - self.lo, offset=4 * self.lo)

    def close(self):
        self._quiesce()
        if self.world > 1:
            for r, b in enumerate(self.peer_d):
                if r != self.rank:
                    b.free()
        for b in (self.d, self.prev, self.work, self.seeds, self.smin, self.stats):
            b.free()


class EmulatedSsspPartition(SsspBackend):
    """All P row blocks on ONE device, each with its OWN full-length dist, prev and work: candidates for another
    block's vertices travel as red.min into that block's dist, exactly as between GPUs."""

    def __init__(self, ctx, at_blocks, bounds, source, max_rounds=4096, window_buckets=0.0):
        self.ctx, self.blocks, self.bounds, self.source = ctx, at_blocks, np.asarray(bounds, dtype=np.int64), int(source)
        self.window = float(window_buckets)
        self.n, self.P = int(self.bounds[-1]), len(at_blocks)
        size = lambda k: int(self.bounds[k + 1] - self.bounds[k])
        self.d = [ctx.empty(max(self.n, 1) * 4) for _ in range(self.P)]
        self.prev = [ctx.empty(max(size(k), 1) * 4) for k in range(self.P)]
        self.work = [ctx.empty(ctx.sssp_block_work_bytes(size(k))) for k in range(self.P)]
        self.seeds = ctx.empty((max_rounds + 2) * 8)
        self.smin = ctx.empty((max_rounds + 2) * 8)
        self.stats = ctx.empty(16)
        self.reset()

    def reset(self):
        self.seeds.zero()
        self.smin.fill_u32(0x7F7FFFFF)
        for k, AT in enumerate(self.blocks):
            self.ctx.sssp_block_init(AT, int(self.bounds[k]), self.n, self.source, self.d[k], self.prev[k], self.work[k])

    def seed(self, r):  # all blocks add to / min into the same words: the all-reduce, for free
        for k, AT in enumerate(self.blocks):
            self.ctx.sssp_block_seed(AT, int(self.bounds[k]), self.n, self.d[k], self.prev[k], self.work[k], self.seeds,
                                     self.smin, r)

    def solve(self, r):
        for k, AT in enumerate(self.blocks):
            self.ctx.sssp_block_solve(AT, int(self.bounds[k]), self.n, self.d[k], self.d, k, self.bounds, self.prev[k],
                                      self.work[k], self.smin, r, self.window)

    def reduce_seeds(self, r):
        pass

    def barrier(self, r):
        pass  # one stream: the next seed passes are ordered after every block's solve

    def read_seeds(self, r):
        return int(self.seeds.download(np.int64, 1, offset=8 * r)[0])

    def stats_total(self):
        r = e = 0
        for k, AT in enumerate(self.blocks):
            self.ctx.sssp_block_finish(AT, int(self.bounds[k]), self.d[k], self.work[k], self.stats)
            w = self.stats.download(np.int64, 2)
            r, e = r + int(w[0]), e + int(w[1])
        return r, e

    def result(self) -> np.ndarray:
        return np.concatenate([self.d[k].download(np.float32, int(self.bounds[k + 1] - self.bounds[k]),
                                                  offset=4 * int(self.bounds[k])) for k in range(self.P)])


def partitioned_sssp(ctx, AT_local, bounds, rank, world, source, *, max_rounds=4096, window_buckets=None, check_every=8,
                     dist=None, torch=None, repeat=1):
    """SSSP distances on a row-partitioned operator.  Returns (backend, stats) with stats = dict(rounds, converged,
    reached, edges_traversed [all ranks], ms [this rank's device time])."""
    be = GpuSsspPartition(ctx, AT_local, bounds, rank, world, source, max_rounds=max_rounds,
                          window_buckets=window_buckets, dist=dist, torch=torch)
    best = None
    for i in range(max(1, repeat)):
        if i:
            be.reset()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        rounds, conv = run_sssp(be, max_rounds, check_every)
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        if best is None or ms < best["ms"]:
            best = {"rounds": rounds, "converged": conv, "ms": ms}
    reached, edges = be.local_stats()
    if world > 1:
        t = torch.tensor([reached, edges], dtype=torch.int64, device=f"cuda:{ctx.device}")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        reached, edges = int(t[0].item()), int(t[1].item())
    best["reached"], best["edges_traversed"] = reached, edges
    return be, best
