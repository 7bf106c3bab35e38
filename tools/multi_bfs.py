#!/usr/bin/env python
"""torchrun entry: row-partitioned BFS on N GPUs (BASELINE.json config 3 family).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/multi_bfs.py --scale 24 --edgefactor 16 [--check]

Every rank generates the same symmetrised RMAT graph on its GPU (bit-identical generator), keeps its
nnz-balanced, word-aligned row block and drops the rest.  Prints one JSON line on rank 0: GTEPS =
(sum of degrees of reached vertices / 2) / device time (max over ranks) -- the single-GPU bench's convention."""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def run(scale, edgefactor, check, seed=0x5EED0024, check_every=8, repeat=3, ctx=None, exchange="peer", trace=False):
    import torch
    import torch.distributed as dist

    import sparseharness_b200 as shb
    from sparseharness_b200 import multi

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    own_ctx = ctx is None
    if own_ctx:
        ctx = shb.Context(local)
        ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    A = ctx.synth_rmat(scale, edgefactor, seed=seed, symmetrise=True)
    info = A.info
    n, nnz = info.n_rows, info.nnz
    rp_head = A.download(0, min(n, 4096))[0]
    root = int(np.argmax(np.diff(rp_head) > 0))
    bounds = multi.word_aligned_bounds(A.balance_rows(world))
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    deg = A.row_lengths()  # symmetric graph: push degree == pull degree, and the row block serves both directions
    single = None
    if check == "sampled":
        # the single-device driver on the whole graph (every rank has it at this point): its levels are tested on
        # sampled row blocks against the BFS local conditions (tools/parity_checks.py), then every rank compares its
        # slice of the partitioned result with it bit for bit.  tests/test_gpu_fullsize.py compares the same
        # single-device levels with the oracle's queue BFS on every vertex.
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import parity_checks as pc
        lv_buf, st1 = ctx.bfs(A, root, AT=A)
        lv_full = lv_buf.download(np.int32, n)
        lv_buf.free()
        begins, cnt = pc.sample_blocks(n, 1 << 14, extra=[int(b) - (1 << 13) for b in bounds[1:-1]] + [lo, hi - (1 << 14)])
        ok = True
        for b in begins:
            rp_s, col_s, _ = A.download(b, cnt)
            ok = ok and pc.bfs_rows_ok(lv_full, rp_s, col_s, b, root)
        single = (lv_full, st1, ok, len(begins) * cnt)
    if world > 1:
        Al = A.slice_rows(lo, hi - lo)
        A.free()
    else:
        Al = A
    be, st = multi.partitioned_bfs(ctx, Al, bounds, rank, world, root, nnz, AT_local=Al, deg=deg, check_every=check_every,
                                   exchange=exchange, dist=dist, torch=torch, repeat=repeat)
    t = torch.tensor([st["ms"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    undirected = st["edges_traversed"] / 2
    out = {"graph": f"RMAT scale {scale} edgefactor {edgefactor} symmetrised+deduplicated", "n": n, "nnz": nnz,
           "n_gpus": world, "exchange": exchange if world > 1 else "none", "schedule": "row-partitioned, direction-optimising; frontier bitmap words / bits stored into every peer by the push and pull kernels",
           "levels": st["levels"], "reached": st["reached"], "ms": ms,
           "timed": "all levels, first launch to the host's last poll; the set-up pass (levels = -1, bitmaps zeroed, isolated "
                    "vertices marked from the bits cached on the operator) is outside -- shb_bfs on one GPU times it too", "gteps": undirected / (ms * 1e6),
           "teps_edges": undirected, "local_rows": hi - lo, "local_nnz": Al.info.nnz,
           "frontier_bytes_per_gpu_per_level": (n + 7) // 8}
    if trace:  # one more traversal, synchronised after every level: per-level device time and direction
        import time
        be.reset()
        per = []
        for lv in range(st["levels"]):
            ctx.synchronize()
            t0 = time.perf_counter()
            be.step(lv)
            be.reduce_count(lv)
            ctx.synchronize()
            mode = int(be.state.download(np.uint32, 1)[0])
            per.append({"level": lv, "us": round((time.perf_counter() - t0) * 1e6, 1), "mode": "pull" if mode else "push",
                        "found": int(be.read_counts(lv, lv + 1)[0])})
        out["trace"] = per
    if check == "sampled":
        lv_full, st1, ok, nrows = single
        got = be.result()
        ok = ok and bool(np.array_equal(got, lv_full[lo:hi])) and st["levels"] == st1.iterations
        ok = ok and st["reached"] == st1.reached and st["edges_traversed"] == st1.edges_traversed
        flag = torch.tensor([int(ok)], dtype=torch.int32, device="cuda")
        if world > 1:
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        out.update(parity_checked=True, parity_ok=bool(flag.item()),
                   parity="levels of the single-device driver satisfy the BFS local conditions on %d sampled rows per rank; "
                          "every rank's slice of the partitioned levels, the level count, reached and traversed-edge "
                          "totals equal the single-device run" % nrows)
    elif check:
        from oracle import oracle as orc  # checker only
        rp, col = orc.gen_rmat(scale, edgefactor, seed=seed, symmetrise=True)
        want, nlev, edges = orc.bfs_levels(rp, col, None, root)
        got = be.result()
        ok = bool(np.array_equal(got, want[lo:hi])) and st["levels"] == nlev and st["reached"] == int((want >= 0).sum())
        ok = ok and st["edges_traversed"] == edges
        flag = torch.tensor([int(ok)], dtype=torch.int32, device="cuda")
        if world > 1:
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        out.update(parity_ok=bool(flag.item()), oracle_levels=nlev, oracle_edges=edges)
    be.close()
    deg.free()
    Al.free()
    return out, rank, world


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=int, default=20)
    ap.add_argument("--edgefactor", type=int, default=16)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--check-every", type=int, default=8)
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"])
    ap.add_argument("--trace", action="store_true")
    a = ap.parse_args()
    out, rank, world = run(a.scale, a.edgefactor, a.check, check_every=a.check_every, exchange=a.exchange, trace=a.trace)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0 if (not a.check or out.get("parity_ok", True)) else 1


if __name__ == "__main__":
    sys.exit(main())
