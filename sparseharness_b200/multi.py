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
