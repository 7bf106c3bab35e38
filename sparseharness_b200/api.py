"""Thin Python view of the C ABI (include/shb.h): Context / Buffer / Csr wrappers over ctypes.

Every compute call goes through libshb.so; nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi
from ._capi import CsrInfo, GraphStats, LaunchCfg, lib

PLUS_TIMES, MIN_PLUS, OR_AND, MAX_MIN = 0, 1, 2, 3
VARIANT_AUTO, VARIANT_SCALAR, VARIANT_VECTOR, VARIANT_STREAM, VARIANT_BLOCKED = 0, 1, 2, 3, 4
TERM_ABSDIFF, TERM_EQUAL = 0, 1
F_NO_CACHE_HINTS, F_NO_Y_SKIP, F_REVERSED_ORDER, F_Y_IS_ONE, F_NO_PANELS = 1, 2, 4, 8, 16
F_STREAM_MERGE, F_STREAM_NO_MERGE, F_STREAM_SCAN, F_STREAM_TMA = 32, 64, 128, 256
_DTYPE = {PLUS_TIMES: np.float32, MIN_PLUS: np.float32, OR_AND: np.int32, MAX_MIN: np.int32}


class ShbError(RuntimeError):
    def __init__(self, status: int):
        self.status = status
        super().__init__(f"{lib.shb_status_string(status).decode()}: {lib.shb_last_error().decode()}")


def _ck(status: int) -> None:
    if status != 0:
        raise ShbError(status)


def device_count() -> int:
    n = C.c_int(0)
    _ck(lib.shb_device_count(C.byref(n)))
    return n.value


def launch_cfg(variant=VARIANT_AUTO, lanes=0, rows_per_group=0, block_threads=0, grid_blocks=0, flags=0):
    return LaunchCfg(variant, lanes, rows_per_group, block_threads, grid_blocks, flags)


def _scalar(semiring, v):
    # an unknown semiring id is rejected by the library (SHB_ERR_UNSUPPORTED), not here
    return np.array([v], dtype=_DTYPE.get(semiring, np.float32))


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Buffer:
    def __init__(self, ctx: "Context", handle, nbytes: int):
        self.ctx, self.h, self.nbytes = ctx, handle, nbytes

    def upload(self, arr: np.ndarray, offset: int = 0) -> int:
        arr = np.ascontiguousarray(arr)
        ns = C.c_uint64(0)
        _ck(lib.shb_buf_upload(self.ctx.h, self.h, offset, _ptr(arr), arr.nbytes, C.byref(ns)))
        return ns.value

    def download(self, dtype, count: int | None = None, offset: int = 0) -> np.ndarray:
        dtype = np.dtype(dtype)
        count = (self.nbytes - offset) // dtype.itemsize if count is None else count
        out = np.empty(count, dtype=dtype)
        _ck(lib.shb_buf_download(self.ctx.h, self.h, offset, _ptr(out), out.nbytes, None))
        return out

    def zero(self):
        _ck(lib.shb_buf_fill_zero(self.ctx.h, self.h, 0, self.nbytes, None))

    def fill_u32(self, pattern: int, n: int | None = None):
        _ck(lib.shb_buf_fill_u32(self.ctx.h, self.h, 0, self.nbytes // 4 if n is None else n, pattern))

    @property
    def device_ptr(self) -> int:
        p = C.c_void_p()
        _ck(lib.shb_buf_device_ptr(self.h, C.byref(p)))
        return p.value

    def ipc_handle(self) -> bytes:
        buf = C.create_string_buffer(64)
        _ck(lib.shb_ipc_export(self.h, buf))
        return buf.raw

    def free(self):
        if self.h:
            lib.shb_buf_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Csr:
    def __init__(self, ctx: "Context", handle):
        self.ctx, self.h = ctx, handle

    @property
    def info(self) -> CsrInfo:
        i = CsrInfo()
        _ck(lib.shb_csr_get_info(self.h, C.byref(i)))
        return i

    def download(self, row_begin=0, row_count=None, dtype=np.float32):
        inf = self.info
        row_count = inf.n_rows - row_begin if row_count is None else row_count
        rp = np.empty(row_count + 1, dtype=np.int64)
        _ck(lib.shb_csr_download(self.ctx.h, self.h, row_begin, row_count, _ptr(rp), None, None))
        nnz = int(rp[-1])
        col = np.empty(nnz, dtype=np.int32)
        val = np.empty(nnz, dtype=dtype) if inf.has_values else None
        _ck(lib.shb_csr_download(self.ctx.h, self.h, row_begin, row_count, _ptr(rp), _ptr(col), _ptr(val)))
        return rp, col, val

    def build_panels(self, n_panels: int = 0) -> int:
        """Column panels for operators whose x exceeds the L2 (shb_csr_build_panels); returns the number built."""
        n = C.c_int(0)
        _ck(lib.shb_csr_build_panels(self.ctx.h, self.h, n_panels, C.byref(n)))
        return n.value

    def transpose(self) -> "Csr":
        h = C.c_void_p()
        _ck(lib.shb_csr_transpose(self.ctx.h, self.h, C.byref(h)))
        return Csr(self.ctx, h)

    def row_lengths(self) -> "Buffer":
        """int32 length of every row, on the device (the push degrees of the partitioned BFS heuristic)."""
        b = self.ctx.empty(max(self.info.n_rows, 1) * 4)
        _ck(lib.shb_csr_row_lengths(self.ctx.h, self.h, b.h))
        return b

    def balance_rows(self, parts: int) -> np.ndarray:
        """Row bounds of `parts` contiguous blocks balanced by non-zeros."""
        b = np.zeros(parts + 1, dtype=np.int64)
        _ck(lib.shb_csr_balance_rows(self.ctx.h, self.h, parts, _ptr(b)))
        return b

    def slice_rows(self, row_begin: int, row_count: int) -> "Csr":
        h = C.c_void_p()
        _ck(lib.shb_csr_slice_rows(self.ctx.h, self.h, row_begin, row_count, C.byref(h)))
        return Csr(self.ctx, h)

    def build_blocked(self, rows_per_consumer: int = 0, make_default: bool = True) -> dict:
        """Build the propagation-blocked copy that VARIANT_BLOCKED runs on (pb.cuh); returns its geometry."""
        _ck(lib.shb_csr_build_blocked(self.ctx.h, self.h, rows_per_consumer, int(make_default)))
        geom = (C.c_int32 * 8)()
        entries, nbytes = C.c_int64(), C.c_uint64()
        _ck(lib.shb_csr_blocked_info(self.h, geom, C.byref(entries), C.byref(nbytes)))
        return {"n_prod": geom[0], "n_pass": geom[1], "n_panel": geom[2], "W": geom[3], "RS": geom[4],
                "cbuf_cap": geom[5], "slot_cap": geom[6], "depth": geom[7], "entries": entries.value, "device_bytes": nbytes.value}

    def free(self):
        if self.h:
            lib.shb_csr_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Context:
    """One device + one in-order stream (the reference Harness's context/queue, inc/harness.h:19-81)."""

    def __init__(self, device: int = 0):
        h = C.c_void_p()
        _ck(lib.shb_ctx_create(device, C.byref(h)))
        self.h = h
        self.device = device

    @property
    def device_name(self) -> str:
        b = C.create_string_buffer(256)
        _ck(lib.shb_ctx_device_name(self.h, b, 256))
        return b.value.decode()

    def set_stream(self, cuda_stream: int):
        _ck(lib.shb_ctx_set_stream(self.h, C.c_void_p(cuda_stream)))

    def synchronize(self):
        _ck(lib.shb_ctx_synchronize(self.h))

    # ---- buffers
    def empty(self, nbytes: int) -> Buffer:
        h = C.c_void_p()
        _ck(lib.shb_buf_create(self.h, nbytes, C.byref(h)))
        return Buffer(self, h, nbytes)

    def csr_from_coo(self, n_rows, n_cols, col_of_entry, row_of_entry, vals=None, truncate=True) -> "Csr":
        """COO tuples -> CSR on the device (shb_csr_from_coo): row = the tuple's second index, tuple order kept in a row."""
        ci = np.ascontiguousarray(col_of_entry, dtype=np.int32)
        rj = np.ascontiguousarray(row_of_entry, dtype=np.int32)
        kind = 0
        if vals is not None:
            vals = np.ascontiguousarray(vals)
            assert vals.dtype in (np.float32, np.int32)
            kind = 0 if vals.dtype == np.float32 else 1
        h = C.c_void_p()
        _ck(lib.shb_csr_from_coo(self.h, n_rows, n_cols, len(ci), _ptr(ci), _ptr(rj), _ptr(vals), kind, int(truncate), C.byref(h), None))
        return Csr(self, h)

    def array(self, arr: np.ndarray) -> Buffer:
        arr = np.ascontiguousarray(arr)
        b = self.empty(max(arr.nbytes, 4))
        if arr.nbytes:
            b.upload(arr)
        return b

    def wrap(self, device_ptr: int, nbytes: int) -> Buffer:
        h = C.c_void_p()
        _ck(lib.shb_buf_wrap(self.h, C.c_void_p(device_ptr), nbytes, C.byref(h)))
        return Buffer(self, h, nbytes)

    def ipc_open(self, handle: bytes, nbytes: int) -> Buffer:
        h = C.c_void_p()
        _ck(lib.shb_ipc_open(self.h, handle, nbytes, C.byref(h)))
        return Buffer(self, h, nbytes)

    # ---- operators
    def csr(self, n_rows, n_cols, row_ptr, col, val=None) -> Csr:
        row_ptr = np.ascontiguousarray(row_ptr, dtype=np.int64)
        col = np.ascontiguousarray(col, dtype=np.int32)
        if val is not None:
            val = np.ascontiguousarray(val)
            assert val.dtype.itemsize == 4
        h = C.c_void_p()
        _ck(lib.shb_csr_create(self.h, n_rows, n_cols, len(col), _ptr(row_ptr), _ptr(col), _ptr(val), 4, C.byref(h), None))
        return Csr(self, h)

    def synth_uniform(self, n_rows, n_cols, k=16, seed=0x5EED0002, val_mode=0, row_begin=0, row_count=None) -> Csr:
        row_count = n_rows - row_begin if row_count is None else row_count
        h = C.c_void_p()
        _ck(lib.shb_synth_uniform(self.h, n_rows, n_cols, k, seed, val_mode, row_begin, row_count, C.byref(h)))
        return Csr(self, h)

    def synth_vector(self, n, seed=0x5EED0003, mode=0) -> Buffer:
        b = self.empty(max(n, 1) * 4)
        _ck(lib.shb_synth_vector(self.h, b.h, n, seed, mode))
        return b

    def synth_rmat(self, scale, edgefactor=16, seed=0x5EED0024, symmetrise=True, value_mode=0, damping=0.85) -> Csr:
        h = C.c_void_p()
        _ck(lib.shb_synth_rmat(self.h, scale, (1 << scale) * edgefactor, seed, int(symmetrise), value_mode, damping,
                               C.byref(h)))
        return Csr(self, h)

    def synth_grid2d(self, w, h_, seed=0x5EED0004) -> Csr:
        h = C.c_void_p()
        _ck(lib.shb_synth_grid2d(self.h, w, h_, seed, C.byref(h)))
        return Csr(self, h)

    # ---- the hot path
    def spmv(self, A: Csr, semiring, x: Buffer, y: Buffer | None, alpha, beta, out: Buffer, cfg: LaunchCfg | None = None,
             timed: bool = True) -> int:
        a, b = _scalar(semiring, alpha), _scalar(semiring, beta)
        ns = C.c_uint64(0)
        _ck(lib.shb_spmv(self.h, A.h, semiring, C.byref(cfg) if cfg else None, x.h, y.h if y else None, _ptr(a), _ptr(b),
                         out.h, C.byref(ns) if timed else None))
        return ns.value

    def spmv_stream_static(self, A: Csr, semiring, x: Buffer, y: Buffer | None, alpha, beta, out: Buffer, timed: bool = True) -> int:
        """shb_spmv_stream_static: the stream schedule with its scratch in the library's static device array (no allocation)."""
        a, b = _scalar(semiring, alpha), _scalar(semiring, beta)
        ns = C.c_uint64(0)
        _ck(lib.shb_spmv_stream_static(self.h, A.h, semiring, x.h, y.h if y else None, _ptr(a), _ptr(b), out.h,
                                       C.byref(ns) if timed else None))
        return ns.value

    def spmv_host(self, A: Csr, semiring, x: np.ndarray, y: np.ndarray | None, alpha, beta, out: np.ndarray,
                  cfg: LaunchCfg | None = None):
        a, b = _scalar(semiring, alpha), _scalar(semiring, beta)
        tot, ker = C.c_uint64(0), C.c_uint64(0)
        _ck(lib.shb_spmv_host(self.h, A.h, semiring, C.byref(cfg) if cfg else None, _ptr(x), _ptr(y), _ptr(a), _ptr(b),
                              _ptr(out), C.byref(tot), C.byref(ker)))
        return tot.value, ker.value

    def spmv_host_batch(self, A: Csr, semiring, xs: list[np.ndarray], ys: list[np.ndarray] | None, alpha, beta,
                        outs: list[np.ndarray], cfg: LaunchCfg | None = None) -> int:
        """Pipelined batch of host vectors (upload i+1 / SpMV i / download i-1 overlap); returns total device ns."""
        a, b = _scalar(semiring, alpha), _scalar(semiring, beta)
        n = len(xs)
        xa = (C.c_void_p * n)(*[x.ctypes.data for x in xs])
        oa = (C.c_void_p * n)(*[o.ctypes.data for o in outs])
        ya = (C.c_void_p * n)(*[y.ctypes.data for y in ys]) if ys is not None else None
        tot = C.c_uint64(0)
        _ck(lib.shb_spmv_host_batch(self.h, A.h, semiring, C.byref(cfg) if cfg else None, n, xa, ya, _ptr(a), _ptr(b), oa,
                                    C.byref(tot)))
        return tot.value

    def spmv_scatter(self, A: Csr, semiring, x: Buffer, y: Buffer | None, alpha, beta, dst: list[Buffer], row_offset,
                     delta=0.0, term=TERM_ABSDIFF, changed: Buffer | None = None, cfg: LaunchCfg | None = None,
                     timed=False, prev_changed: Buffer | None = None) -> int:
        a, b = _scalar(semiring, alpha), _scalar(semiring, beta)
        arr = (C.c_void_p * len(dst))(*[d.h for d in dst])
        ns = C.c_uint64(0)
        _ck(lib.shb_spmv_scatter(self.h, A.h, semiring, C.byref(cfg) if cfg else None, x.h, y.h if y else None, _ptr(a),
                                 _ptr(b), arr, len(dst), row_offset, delta, term, changed.h if changed else None,
                                 prev_changed.h if prev_changed else None, C.byref(ns) if timed else None))
        return ns.value

    def bfs_block_state_bytes(self) -> int:
        b = C.c_uint64(0)
        _ck(lib.shb_bfs_block_state_bytes(C.byref(b)))
        return b.value

    def bfs_block_counter_bytes(self) -> tuple[int, int]:
        """(bytes of the counters buffer, number of frontier slots in its ring)."""
        b, r = C.c_uint64(0), C.c_uint32(0)
        _ck(lib.shb_bfs_block_counter_bytes(C.byref(b), C.byref(r)))
        return b.value, r.value

    def bfs_block_init(self, A: Csr, AT: Csr | None, row_offset, n_total, nnz_total, root, deg: Buffer | None,
                       levels: Buffer, bm0: Buffer, bm1: Buffer, bm2: Buffer, visited: Buffer, counters: Buffer,
                       state: Buffer, staged: bool = False):
        _ck(lib.shb_bfs_block_init(self.h, A.h, AT.h if AT else None, row_offset, n_total, nnz_total, root,
                                   deg.h if deg else None, levels.h, bm0.h, bm1.h, bm2.h, visited.h, counters.h, state.h,
                                   int(staged)))

    def bfs_block_publish(self, slot, counters: Buffer, state: Buffer, peer_counters: list[Buffer], self_index: int,
                          barrier_target: int):
        arr = (C.c_void_p * len(peer_counters))(*[d.h for d in peer_counters])
        _ck(lib.shb_bfs_block_publish(self.h, slot, counters.h, state.h, arr, len(peer_counters), self_index,
                                      barrier_target))

    def bfs_block_level(self, A: Csr, AT: Csr | None, row_offset, n_total, level, deg: Buffer | None, frontier: Buffer,
                        nxt: list[Buffer], self_index: int, spare: Buffer, visited: Buffer, levels: Buffer,
                        counters: Buffer, state: Buffer, peer_counters: list[Buffer] | None = None,
                        barrier_target: int = 0):
        arr = (C.c_void_p * len(nxt))(*[d.h for d in nxt])
        pc = (C.c_void_p * len(peer_counters))(*[d.h for d in peer_counters]) if peer_counters else None
        _ck(lib.shb_bfs_block_level(self.h, A.h, AT.h if AT else None, row_offset, n_total, level, deg.h if deg else None,
                                    frontier.h, arr, len(nxt), self_index, spare.h, visited.h, levels.h, counters.h,
                                    state.h, pc, barrier_target))

    def bfs_block_label_frontier(self, A: Csr, row_offset, n_total, level, frontier: Buffer, spare: Buffer, visited: Buffer,
                                 levels: Buffer, counters: Buffer, state: Buffer):
        _ck(lib.shb_bfs_block_label_frontier(self.h, A.h, row_offset, n_total, level, frontier.h, spare.h, visited.h, levels.h,
                                             counters.h, state.h))

    def bfs_block_finish(self, A: Csr, levels: Buffer, counters: Buffer, state: Buffer):
        _ck(lib.shb_bfs_block_finish(self.h, A.h, levels.h, counters.h, state.h))

    def sssp_block_work_bytes(self, n_local: int) -> int:
        b = C.c_uint64(0)
        _ck(lib.shb_sssp_block_work_bytes(n_local, C.byref(b)))
        return b.value

    def sssp_block_init(self, AT: Csr, row_offset, n_total, source, dist: Buffer, prev: Buffer, work: Buffer):
        _ck(lib.shb_sssp_block_init(self.h, AT.h, row_offset, n_total, source, dist.h, prev.h, work.h))

    def sssp_block_seed(self, AT: Csr, row_offset, n_total, dist: Buffer, prev: Buffer, work: Buffer,
                        seed_count: Buffer, seed_min: Buffer | None, word: int):
        _ck(lib.shb_sssp_block_seed(self.h, AT.h, row_offset, n_total, dist.h, prev.h, work.h, seed_count.h,
                                    seed_min.h if seed_min else None, word))

    def sssp_block_solve(self, AT: Csr, row_offset, n_total, dist: Buffer, peer_dist: list[Buffer], self_index: int,
                         bounds: np.ndarray, prev: Buffer, work: Buffer, seed_min: Buffer | None = None, word: int = 0,
                         window_buckets: float = 0.0, max_phases: int = 0):
        arr = (C.c_void_p * len(peer_dist))(*[d.h for d in peer_dist])
        b = np.ascontiguousarray(bounds, dtype=np.int64)
        _ck(lib.shb_sssp_block_solve(self.h, AT.h, row_offset, n_total, dist.h, arr, len(peer_dist), self_index, _ptr(b),
                                     prev.h, work.h, seed_min.h if seed_min else None, word, window_buckets, max_phases))

    def sssp_block_finish(self, A_rows: Csr, row_offset, dist: Buffer, work: Buffer, stats_out: Buffer):
        _ck(lib.shb_sssp_block_finish(self.h, A_rows.h, row_offset, dist.h, work.h, stats_out.h))

    def iterate(self, A: Csr, semiring, x: Buffer, y0: Buffer | None, alpha, beta, work: Buffer, delta=1e-4, term=None,
                y_rebind=True, max_iters=1000, check_every=16, per_iter=False, cfg: LaunchCfg | None = None):
        """Returns (result_buffer, n_iters, total_ns, per_iter_ns|None, converged)."""
        term = (TERM_ABSDIFF if semiring in (PLUS_TIMES, MIN_PLUS) else TERM_EQUAL) if term is None else term
        a, b = _scalar(semiring, alpha), _scalar(semiring, beta)
        n_it, in_work, tot = C.c_uint32(0), C.c_int(0), C.c_uint64(0)
        per = np.zeros(max_iters, dtype=np.uint64) if per_iter else None
        st = lib.shb_iterate(self.h, A.h, semiring, C.byref(cfg) if cfg else None, x.h, y0.h if y0 else None, _ptr(a),
                             _ptr(b), work.h, delta, term, int(y_rebind), max_iters, check_every,
                             per.ctypes.data_as(_capi.u64p) if per is not None else None, C.byref(n_it),
                             C.byref(in_work), C.byref(tot))
        if st not in (0, 7):
            raise ShbError(st)
        res = work if in_work.value else x
        return res, n_it.value, tot.value, (per[: n_it.value] if per is not None else None), st == 0

    def bfs(self, A: Csr, root=0, AT: Csr | None = None, max_levels=0):
        levels = self.empty(A.info.n_rows * 4)
        st = GraphStats()
        _ck(lib.shb_bfs(self.h, A.h, AT.h if AT else None, root, levels.h, max_levels, C.byref(st)))
        return levels, st

    def sssp(self, A: Csr, source=0, AT: Csr | None = None, max_rounds=0):
        dist = self.empty(A.info.n_rows * 4)
        st = GraphStats()
        _ck(lib.shb_sssp(self.h, A.h, AT.h if AT else None, source, dist.h, max_rounds, C.byref(st)))
        return dist, st

    def pagerank(self, A: Csr, damping=0.85, delta=1e-4, max_iters=1000, check_every=8, cfg: LaunchCfg | None = None,
                 both=False):
        """Returns (x_k, stats); with both=True (x_k, x_{k-1}, stats): the two ping-pong buffers."""
        n = A.info.n_rows
        x, work = self.empty(n * 4), self.empty(n * 4)
        x_init = np.float32(1.0) / np.float32(n)
        beta = (np.float32(1.0) - np.float32(damping)) / np.float32(n)
        in_work = C.c_int(0)
        st = GraphStats()
        s = lib.shb_pagerank(self.h, A.h, C.byref(cfg) if cfg else None, x.h, work.h, C.c_float(x_init), C.c_float(beta),
                             delta, max_iters, check_every, C.byref(in_work), C.byref(st))
        if s not in (0, 7):
            raise ShbError(s)
        if both:
            return (work, x, st) if in_work.value else (x, work, st)
        (x if in_work.value else work).free()
        return (work if in_work.value else x), st

    def power_iteration(self, A: Csr, x0: np.ndarray, delta=1e-4, max_iters=10, cfg: LaunchCfg | None = None):
        """Returns (x_k as numpy, eigenvalue estimate, stats)."""
        n = A.info.n_rows
        x, work = self.array(np.ascontiguousarray(x0, dtype=np.float32)), self.empty(max(n, 1) * 4)
        in_work, ev, st = C.c_int(0), C.c_float(0), GraphStats()
        s = lib.shb_power_iteration(self.h, A.h, C.byref(cfg) if cfg else None, x.h, work.h, delta, max_iters, C.byref(ev), C.byref(in_work),
                                    C.byref(st))
        if s not in (0, 7):
            raise ShbError(s)
        res = (work if in_work.value else x).download(np.float32, n)
        x.free(); work.free()
        return res, ev.value, st

    def close(self):
        if self.h:
            lib.shb_ctx_destroy(self.h)
            self.h = None


class _BorrowedCsr(Csr):
    """A block owned by a GroupPartition: a view for downloads, never destroyed from Python."""

    def free(self):
        self.h = None


class _BorrowedContext(Context):
    """A context owned by a Group (not destroyed by close())."""

    def __init__(self, handle, device):  # noqa: super().__init__ deliberately not called
        self.h, self.device = handle, device

    def close(self):
        self.h = None


class GroupPartition:
    """An operator row-partitioned over a Group (shb_gpart): block i resident on device i."""

    def __init__(self, group: "Group", handle):
        self.group, self.h = group, handle

    @property
    def info(self):
        n_rows, n_cols, nnz = C.c_int64(), C.c_int64(), C.c_int64()
        bounds = np.zeros(self.group.n + 1, dtype=np.int64)
        _ck(lib.shb_gpart_info(self.h, C.byref(n_rows), C.byref(n_cols), C.byref(nnz), _ptr(bounds)))
        return {"n_rows": n_rows.value, "n_cols": n_cols.value, "nnz": nnz.value, "bounds": bounds}

    def block(self, i: int) -> Csr:
        """Borrowed view of block i (for downloads / parity checks); freed with the partition."""
        a = C.c_void_p()
        _ck(lib.shb_gpart_block(self.h, i, C.byref(a), None))
        return _BorrowedCsr(self.group.ctx(i), a)

    def free(self):
        if self.h:
            _ck(lib.shb_gpart_destroy(self.h))
            self.h = None


class Group:
    """N devices of one box driven by ONE host thread through shb_group_* (include/shb.h): contexts, peer access and
    the round / level loops live in libshb.so; no torch.distributed, no second process."""

    def __init__(self, n_dev: int, devices=None):
        h = C.c_void_p()
        dv = None if devices is None else (C.c_int * n_dev)(*devices)
        _ck(lib.shb_group_create(n_dev, dv, C.byref(h)))
        self.h, self.n = h, n_dev
        self._ctx = {}

    def ctx(self, i: int) -> Context:
        if i not in self._ctx:
            c = C.c_void_p()
            _ck(lib.shb_group_ctx(self.h, i, C.byref(c)))
            d = C.c_int()
            _ck(lib.shb_ctx_device(c, C.byref(d)))
            self._ctx[i] = _BorrowedContext(c, d.value)
        return self._ctx[i]

    def partition_host(self, n_rows, n_cols, row_ptr, col, val=None, symmetric=False) -> GroupPartition:
        row_ptr = np.ascontiguousarray(row_ptr, dtype=np.int64)
        col = np.ascontiguousarray(col, dtype=np.int32)
        if val is not None:
            val = np.ascontiguousarray(val)
            assert val.dtype.itemsize == 4
        h = C.c_void_p()
        _ck(lib.shb_group_partition_host(self.h, n_rows, n_cols, _ptr(row_ptr), _ptr(col), _ptr(val), int(symmetric), C.byref(h)))
        return GroupPartition(self, h)

    def adopt(self, blocks, bounds, n_cols, at_blocks=None) -> GroupPartition:
        """blocks[i]: a Csr built on self.ctx(i); ownership moves to the partition."""
        arr = (C.c_void_p * self.n)(*[b.h for b in blocks])
        at = None if at_blocks is None else (C.c_void_p * self.n)(*[b.h for b in at_blocks])
        bounds = np.ascontiguousarray(bounds, dtype=np.int64)
        h = C.c_void_p()
        _ck(lib.shb_group_adopt(self.h, arr, at, _ptr(bounds), n_cols, C.byref(h)))
        for b in list(blocks) + list(at_blocks or []):
            b.h = None
        return GroupPartition(self, h)

    def iterate(self, part: GroupPartition, semiring, x0, y0, alpha, beta, delta=1e-4, term=None, y_mode=1, max_iters=1000,
                check_every=16, cfg: LaunchCfg | None = None, want_prev=False):
        """Returns (result, prev|None, n_iters, total_ns, converged)."""
        dt = _DTYPE[semiring]
        n = part.info["n_rows"]
        term = (TERM_ABSDIFF if semiring in (PLUS_TIMES, MIN_PLUS) else TERM_EQUAL) if term is None else term
        a, b = _scalar(semiring, alpha), _scalar(semiring, beta)
        x0 = np.ascontiguousarray(x0, dtype=dt)
        y0 = None if y0 is None else np.ascontiguousarray(y0, dtype=dt)
        res = np.empty(n, dtype=dt)
        prev = np.empty(n, dtype=dt) if want_prev else None
        it, ns = C.c_uint32(0), C.c_uint64(0)
        st = lib.shb_group_iterate(self.h, part.h, semiring, C.byref(cfg) if cfg else None, _ptr(x0), 0, _ptr(y0), _ptr(a), _ptr(b),
                                   delta, term, y_mode, max_iters, check_every, _ptr(res), _ptr(prev), C.byref(it), C.byref(ns))
        if st not in (0, 7):
            raise ShbError(st)
        return res, prev, it.value, ns.value, st == 0

    def pagerank(self, part: GroupPartition, damping=0.85, delta=1e-4, max_iters=1000, check_every=8, cfg: LaunchCfg | None = None,
                 want_result=True, want_prev=False):
        n = part.info["n_rows"]
        x_init = np.float32(1.0) / np.float32(n)
        beta = (np.float32(1.0) - np.float32(damping)) / np.float32(n)
        res = np.empty(n, dtype=np.float32) if want_result else None
        prev = np.empty(n, dtype=np.float32) if want_prev else None
        st = GraphStats()
        s = lib.shb_group_pagerank(self.h, part.h, C.byref(cfg) if cfg else None, C.c_float(x_init), C.c_float(beta), delta, max_iters,
                                   check_every, _ptr(res), _ptr(prev), C.byref(st))
        if s not in (0, 7):
            raise ShbError(s)
        return res, prev, st

    def bfs(self, part: GroupPartition, root=0, max_levels=0, check_every=8, want_levels=True):
        n = part.info["n_rows"]
        lv = np.empty(n, dtype=np.int32) if want_levels else None
        st = GraphStats()
        _ck(lib.shb_group_bfs(self.h, part.h, root, _ptr(lv), max_levels, check_every, C.byref(st)))
        return lv, st

    def sssp(self, part: GroupPartition, source=0, window_buckets=-1.0, max_rounds=0, check_every=8, want_dist=True):
        n = part.info["n_rows"]
        d = np.empty(n, dtype=np.float32) if want_dist else None
        st = GraphStats()
        _ck(lib.shb_group_sssp(self.h, part.h, source, _ptr(d), C.c_float(window_buckets), max_rounds, check_every, C.byref(st)))
        return d, st

    def close(self):
        if self.h:
            _ck(lib.shb_group_destroy(self.h))
            self.h = None
