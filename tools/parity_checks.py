"""Bounded-cost parity gates for the full-size graph runs (bench.py, tools/multi_*.py, tests).

CHECKER CODE: uses the CPU oracle (oracle/), so it is imported only by bench.py's parity legs, the tools' --check
paths and tests/ -- never by the product package.  Each check looks at sampled row blocks of the operator that
actually ran (downloaded from the device) and at the full result vector, and tests the local condition whose
satisfaction on EVERY row characterises the exact answer:

  BFS levels   level[root] = 0; a reached row has a neighbour exactly one level below and none further below;
               an unreached row (-1) has no reached neighbour          (unique solution = queue-BFS levels)
  SSSP         dist[source] = 0; dist[r] = min_k fl(dist[col k] + val k) elsewhere, in float arithmetic
               (the Bellman fixed point of example/sssp/kernel5.json:3's recurrence; unique for positive weights)
  PageRank     x_new[r] = (sum_k x_prev[col k] * val k) * 1 + 1 * beta with the oracle's own loop
               (example/pr/kernel5.json:3, canonical mode Q9), 1e-5 relative in the max norm of the sample

The full-size, every-row comparisons against the oracle are in tests/test_gpu_fullsize.py; these samples are what
bench.py can afford inside every run, at every GPU count.
"""
from __future__ import annotations

import numpy as np


def sample_blocks(n_rows: int, count: int = 4096, extra=()):
    """[begin, ...] of row blocks: first, middle, last + any `extra` begins (clamped)."""
    count = min(count, n_rows)
    begins = {0, max(0, n_rows // 2 - count // 2), max(0, n_rows - count)}
    for b in extra:
        begins.add(int(min(max(0, b), max(0, n_rows - count))))
    return sorted(begins), count


def bfs_rows_ok(levels_full: np.ndarray, rp: np.ndarray, col: np.ndarray, row_begin: int, root: int) -> bool:
    """Rows [row_begin, row_begin + len(rp) - 1) of the pull CSR against the full level vector."""
    n = len(rp) - 1
    if n == 0:
        return True
    lv = levels_full[row_begin:row_begin + n].astype(np.int64)
    big = np.int64(1) << 40
    nb = levels_full[col].astype(np.int64)
    nb = np.where(nb < 0, big, nb)
    lens = np.diff(rp)
    rowmin = np.full(n, big, np.int64)
    nz = lens > 0
    if len(col):
        red = np.minimum.reduceat(nb, rp[:-1][nz])
        rowmin[nz] = red
    rows = np.arange(row_begin, row_begin + n)
    want = np.where(rowmin >= big, -1, rowmin + 1)
    want = np.where(rows == root, 0, want)
    # a row whose only way in is through itself (self loop) cannot be reached: rowmin then includes its own level,
    # which is >= want, so the formula above still holds
    return bool(np.array_equal(np.where(rows == root, 0, lv), want))


def sssp_rows_ok(dist_full: np.ndarray, rp: np.ndarray, col: np.ndarray, val: np.ndarray, row_begin: int,
                 source: int) -> bool:
    n = len(rp) - 1
    if n == 0:
        return True
    d = dist_full[row_begin:row_begin + n]
    with np.errstate(over="ignore"):
        cand = (np.abs(dist_full[col]) + np.abs(val.astype(np.float32))).astype(np.float32)  # absadd, one rounded add
    flt_max = np.float32(3.4028235e38)
    cand = np.where(np.isfinite(cand), cand, flt_max)  # FLT_MAX + w stays FLT_MAX for |w| < 2^103
    lens = np.diff(rp)
    rowmin = np.full(n, flt_max, np.float32)
    nz = lens > 0
    if len(col):
        rowmin[nz] = np.minimum.reduceat(cand, rp[:-1][nz])
    rows = np.arange(row_begin, row_begin + n)
    want = np.where(rows == source, np.float32(0), rowmin)
    return bool(np.array_equal(want.view(np.uint32), d.view(np.uint32)))


def pagerank_rows_err(orc, x_prev_full: np.ndarray, x_new_rows: np.ndarray, rp, col, val, beta: float) -> float:
    n = len(rp) - 1
    if n == 0:
        return 0.0
    want = orc.spmv(orc.PLUS_TIMES, rp, col, val, x_prev_full, np.ones(n, np.float32), 1.0, beta, order=0,
                    threads=orc.num_threads())
    scale = float(np.abs(want).max())
    return float(np.abs(x_new_rows.astype(np.float64) - want.astype(np.float64)).max() / max(scale, 1e-300))


def checksum64(a: np.ndarray) -> int:
    """Order-independent 64-bit checksum of a 4-byte vector's bit patterns (sum mod 2^64 of word * odd weights is
    order DEPENDENT; here: plain sum of the words plus sum of their squares, both mod 2^64)."""
    w = np.ascontiguousarray(a).view(np.uint32).astype(np.uint64)
    return int((w.sum(dtype=np.uint64) + (w * w).sum(dtype=np.uint64)) & np.uint64(0xFFFFFFFFFFFFFFFF))
