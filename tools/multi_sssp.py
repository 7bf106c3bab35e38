#!/usr/bin/env python
"""torchrun entry: row-partitioned SSSP on N GPUs (BASELINE.json config 4 family: 2-D grid, min-plus).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/multi_sssp.py --width 8192 --height 8192 [--check]

Every rank generates the same grid on its GPU (bit-identical generator) and keeps its nnz-balanced row block (a
slab of grid rows).  Prints one JSON line on rank 0: GTEPS = sum of degrees of reached vertices / device time (max
over ranks) -- the single-GPU bench's convention."""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def run(width, height, check, seed=0x5EED0004, source=0, repeat=2, ctx=None, window=None, check_every=8):
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
    if ctx is None:
        ctx = shb.Context(local)
        ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    G = ctx.synth_grid2d(width, height, seed=seed)  # symmetric: the rows of a block are its out-edges too
    info = G.info
    n, nnz = info.n_rows, info.nnz
    bounds = G.balance_rows(world)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    single = None
    if check == "sampled":
        # see tools/multi_bfs.py: single-device distances tested on sampled rows (either side of every block boundary)
        # against the Bellman fixed-point condition, then compared slice by slice with the partitioned result
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import parity_checks as pc
        d_buf, st1 = ctx.sssp(G, source, AT=G)
        d_full = d_buf.download(np.float32, n)
        d_buf.free()
        begins, cnt = pc.sample_blocks(n, 1 << 14, extra=[int(b) - (1 << 13) for b in bounds[1:-1]] + [lo, hi - (1 << 14)])
        ok = bool(st1.converged)
        for b in begins:
            rp_s, col_s, val_s = G.download(b, cnt)
            ok = ok and pc.sssp_rows_ok(d_full, rp_s, col_s, val_s, b, source)
        single = (d_full, st1, ok, len(begins) * cnt, pc.checksum64(d_full))
    if world > 1:
        Gl = G.slice_rows(lo, hi - lo)
        G.free()
    else:
        Gl = G
    be, st = multi.partitioned_sssp(ctx, Gl, bounds, rank, world, source, window_buckets=window, check_every=check_every, dist=dist, torch=torch,
                                    repeat=repeat)
    t = torch.tensor([st["ms"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    out = {"graph": f"2-D grid {width}x{height} 4-neighbour, integer weights 1..255", "n": n, "nnz": nnz, "n_gpus": world,
           "schedule": "row-partitioned block-asynchronous near-far; boundary candidates as red.min into the owner's dist "
                       "over peer mappings, one 8-byte all-reduce per round",
           "window_buckets": be.window, "rounds": st["rounds"], "converged": bool(st["converged"]), "reached": st["reached"], "ms": ms,
           "gteps": st["edges_traversed"] / (ms * 1e6), "local_rows": hi - lo, "local_nnz": Gl.info.nnz}
    if check == "sampled":
        d_full, st1, ok, nrows, csum = single
        ok = ok and bool(np.array_equal(be.result().view(np.uint32), d_full[lo:hi].view(np.uint32))) and bool(st["converged"])
        ok = ok and st["reached"] == st1.reached
        flag = torch.tensor([int(ok)], dtype=torch.int32, device="cuda")
        if world > 1:
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        out.update(parity_checked=True, parity_ok=bool(flag.item()), dist_checksum64=csum,
                   single_device_ms=st1.total_ns / 1e6,
                   parity="distances of the single-device driver satisfy the Bellman fixed-point condition (float adds) on "
                          "%d sampled rows per rank incl. both sides of every block boundary; every rank's slice of the "
                          "partitioned distances equals them bit for bit" % nrows)
    elif check:
        from oracle import oracle as orc  # checker only
        rp, col, val = orc.gen_grid2d(width, height, seed=seed)
        want = orc.sssp(rp, col, val, source)
        ok = bool(np.array_equal(be.result(), want[lo:hi])) and st["converged"] and st["reached"] == int((want < orc.FLT_MAX).sum())
        flag = torch.tensor([int(ok)], dtype=torch.int32, device="cuda")
        if world > 1:
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        out.update(parity_ok=bool(flag.item()))
    be.close()
    Gl.free()
    return out, rank, world


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=1024)
    ap.add_argument("--height", type=int, default=1024)
    ap.add_argument("--source", type=int, default=0)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--check-every", type=int, default=8)
    ap.add_argument("--window", type=float, default=None, help="distance window of a round in near-far buckets (0 = none)")
    a = ap.parse_args()
    out, rank, world = run(a.width, a.height, a.check, source=a.source, window=a.window, check_every=a.check_every)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0 if (not a.check or out.get("parity_ok", True)) else 1


if __name__ == "__main__":
    sys.exit(main())
