#!/usr/bin/env python
"""Small operators through the schedules the harness executables do not reach: the blocked schedule (shared-memory
scatter / reduce, TMA slice staging, ring counters), column panels, the stream schedule with straddling rows, and the
group drivers on one device.  Run under compute-sanitizer by tools/gpu_sanitize.sh; results are checked by the test
suite, here only the memory / race behaviour matters."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sparseharness_b200 as shb  # noqa: E402

rng = np.random.default_rng(5)
ctx = shb.Context(0)
n, m = 4000, 3500
lens = rng.poisson(9, n)
lens[7] = 9000
rp = np.zeros(n + 1, np.int64)
np.cumsum(lens, out=rp[1:])
col = rng.integers(0, m, rp[-1]).astype(np.int32)
val = rng.integers(-3, 4, rp[-1]).astype(np.float32)
x = ctx.array(rng.integers(-3, 4, m).astype(np.float32))
out = ctx.empty(n * 4)
yb = ctx.array(np.zeros(n, np.float32))
A = ctx.csr(n, m, rp, col, val)
for rows in (0, 7):
    A.build_blocked(rows, make_default=False)
    for sem in (shb.PLUS_TIMES, shb.MIN_PLUS):
        ctx.spmv(A, sem, x, yb, 1.0, 0.0, out, shb.launch_cfg(variant=shb.VARIANT_BLOCKED))
for v in (shb.VARIANT_STREAM, shb.VARIANT_VECTOR, shb.VARIANT_SCALAR):
    ctx.spmv(A, shb.PLUS_TIMES, x, None, 1.0, 0.0, out, shb.launch_cfg(variant=v))
A.build_panels(3)
ctx.spmv(A, shb.PLUS_TIMES, x, None, 1.0, 0.0, out)
ctx.synchronize()
g = shb.Group(1)
k = 1500
lens = rng.poisson(6, k)
rp = np.zeros(k + 1, np.int64)
np.cumsum(lens, out=rp[1:])
col = rng.integers(0, k, rp[-1]).astype(np.int32)
val = (rng.integers(1, 9, rp[-1])).astype(np.float32)
part = g.partition_host(k, k, rp, col, val, symmetric=False)
g.pagerank(part, delta=1e-6, max_iters=30)
g.bfs(part, int(np.argmax(np.bincount(col, minlength=k) > 0)))
g.sssp(part, 0)
part.free()
g.close()
ctx.close()
print("sanitize driver done")
