#!/usr/bin/env python
"""bench.py -- headline benchmark of the semiring CSR SpMV hot path (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--extras 0|1]

A "step" is one SpMV pass (out = A x, float plus-times, alpha 1 beta 0) over the synthetic
uniform-random CSR operator of config 2: 16 777 216 rows x 16 nnz/row per GPU, columns uniform in
[0, 16 777 216) (seeds 0x5EED0002 / 0x5EED0003), generated on the device, resident in HBM
before the timed region.  N > 1 (torchrun, one rank per GPU): every rank owns an independent block of
16M rows of a (N*16M) x 16M operator with x replicated -- no data-path collective, weak scaling.
Metric: effective HBM GB/s = algorithmic bytes / time, algorithmic bytes per SpMV =
nnz*(4+4) + (rows+1)*4 + cols*4 [x once] + rows*4 [out] = 2 348 810 244 B per GPU.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ROWS_PER_GPU = 1 << 24
NNZ_PER_ROW = 16
N_COLS = 1 << 24
ALGO_BYTES = ROWS_PER_GPU * NNZ_PER_ROW * 8 + (ROWS_PER_GPU + 1) * 4 + N_COLS * 4 + ROWS_PER_GPU * 4  # 2 348 810 244
FLOPS = 2 * ROWS_PER_GPU * NNZ_PER_ROW
CPU_SAMPLE_ROWS = 1 << 21  # bounded CPU sample: first 2M rows (32M non-zeros) of the same operator
REF_SAMPLE_ROWS = 1 << 18  # sample handed to the unmodified reference (text file + fscanf loader): 4M non-zeros
WORKLOAD = ("configs[1]: synthetic uniform-random CSR, 16777216 rows x 16 nnz/row per GPU, float plus-times SpMV "
            "(alpha=1, beta=0)")


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
        except Exception:
            pass
    return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


class ClockSampler:
    """Samples SM clock + throttle reasons of one GPU while the timed region runs (NVML, 50 ms period)."""

    def __init__(self, index: int):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # pragma: no cover
            self.nv, self.err = None, repr(e)

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4,
                 "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        if self.nv:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def cpu_sample_inputs(orc):
    rp, col, val = orc.gen_uniform(ROWS_PER_GPU, N_COLS, NNZ_PER_ROW, row_begin=0, row_count=CPU_SAMPLE_ROWS)
    x = orc.gen_vector(N_COLS)
    nbytes = len(col) * 8 + (CPU_SAMPLE_ROWS + 1) * 4 + N_COLS * 4 + CPU_SAMPLE_ROWS * 4
    return rp, col, val, x, nbytes


def time_cpu(orc, rp, col, val, x, threads, steps, warmup):
    y = np.zeros(CPU_SAMPLE_ROWS, np.float32)
    for _ in range(warmup):
        orc.spmv(orc.PLUS_TIMES, rp, col, val, x, y, 1.0, 0.0, order=0, threads=threads)
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        orc.spmv(orc.PLUS_TIMES, rp, col, val, x, y, 1.0, 0.0, order=0, threads=threads)
        times.append(time.perf_counter() - t0)
    return times


def time_reference_gold(orc, warmup, steps):
    """The UNMODIFIED reference CPU path (oracle/_ref/ref_driver: the reference's loader + Gold<float>::spmv,
    inc/spmv_gold.h:9-28, compiled from /root/reference by oracle/Makefile) on the first REF_SAMPLE_ROWS rows of
    the config-2 operator, handed over as the Matrix Market text file its loader needs.  One step = one
    Gold::spmv call (single threaded by construction; it takes the matrix by value).  None if the driver is absent."""
    if not os.path.exists(orc.REF_DRIVER):
        return None
    import tempfile
    rp, col, val = orc.gen_uniform(ROWS_PER_GPU, N_COLS, NNZ_PER_ROW, row_begin=0, row_count=REF_SAMPLE_ROWS)
    nbytes = len(col) * 8 + (REF_SAMPLE_ROWS + 1) * 4 + N_COLS * 4 + REF_SAMPLE_ROWS * 4
    fd, path = tempfile.mkstemp(suffix=".mtx", prefix="shb_ref_sample_")
    os.close(fd)
    try:
        orc.write_mtx(path, REF_SAMPLE_ROWS, N_COLS, rp, col, val, second_index_is_row=True)
        r = orc.reference_gold_time(path, warmup, steps)
    except Exception:
        r = None
    finally:
        os.unlink(path)
    if r is None:
        return None
    mean_ns, best_ns = r
    return {"mean_s": mean_ns / 1e9, "best_s": best_ns / 1e9, "bytes": nbytes, "nnz": len(col),
            "sample": f"first {REF_SAMPLE_ROWS} rows ({len(col)} non-zeros) of the config-2 operator as a Matrix Market file "
                      f"through the reference's own loader and Gold<float>::spmv (oracle/_ref/ref_driver), 1 thread"}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores: oracle/_ref
    (the unmodified loader + Gold<float>::spmv) on a bounded sample of config 2.  Gold is single threaded by
    construction, so cores = 1; the OpenMP restatement on every core is reported beside it ("port_all_cores").
    Falls back to that restatement (kind "port") only if oracle/_ref is missing."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import oracle as orc
    ref = time_reference_gold(orc, max(args.warmup, 1), max(args.steps, 1))
    rp, col, val, x, nbytes = cpu_sample_inputs(orc)
    threads = orc.num_threads()
    port_times = time_cpu(orc, rp, col, val, x, threads, min(args.steps, 20), 1)
    port = {"value": nbytes / float(np.mean(port_times)) / 1e9, "unit": "GB/s", "cores": threads,
            "sample": f"first {CPU_SAMPLE_ROWS} rows ({len(col)} non-zeros), oracle restatement of inc/spmv_gold.h:17-26, OpenMP over rows"}
    if ref is not None:
        mean, gbs, kind, cores, sample, nnz = ref["mean_s"], ref["bytes"] / ref["mean_s"] / 1e9, "reference", 1, ref["sample"], ref["nnz"]
    else:
        mean = float(np.mean(port_times))
        gbs, kind, cores, sample, nnz = port["value"], "port", threads, port["sample"], len(col)
    line = {"impl": "reference", "metric": "spmv_effective_hbm_gbs", "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": mean * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": sample},
            "gflops": 2 * nnz / mean / 1e9,
            "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": kind, "sample": sample,
                             "port_all_cores": port},
            "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)
    return 0


def graph_extras(ctx, shb, orc):
    """GTEPS of the graph drivers on one GPU (configs 3-5 families), reported next to the headline line.  Every
    number is gated by a sampled parity check (tools/parity_checks.py); a mismatch aborts the run."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import parity_checks as pc
    out = {}

    def guarded(name, fn):
        try:
            out[name] = fn()
        except SystemExit:
            raise
        except Exception as e:  # keep the headline line alive
            out[name] = {"error": repr(e)[:300]}

    def bfs():
        A = ctx.synth_rmat(24, 16, seed=0x5EED0024, symmetrise=True)
        info = A.info
        rp = A.download(0, 4096)[0]
        root = int(np.argmax(np.diff(rp) > 0))
        best = None
        for _ in range(3):
            levels, st = ctx.bfs(A, root, AT=A)
            if best is None or st.total_ns < best.total_ns:
                best = st
            lv = levels.download(np.int32, info.n_rows)
            levels.free()
        begins, cnt = pc.sample_blocks(info.n_rows, 1 << 14)
        ok = all(pc.bfs_rows_ok(lv, *A.download(b, cnt)[:2], b, root) for b in begins)
        if not ok:
            raise SystemExit("bench: BFS levels violate the BFS local conditions -- refusing to report a number")
        undirected = best.edges_traversed / 2
        r = {"graph": "RMAT scale 24 edgefactor 16 symmetrised+deduplicated", "n": info.n_rows, "nnz": info.nnz,
             "levels": best.iterations, "reached": best.reached, "ms": best.total_ns / 1e6,
             "gteps": undirected / best.total_ns, "teps_edges": undirected, "parity_checked": True,
             "parity": f"BFS local conditions (parent one level below, none further, unreached closed) on {len(begins) * cnt} sampled rows"}
        A.free()
        return r

    def sssp():
        G = ctx.synth_grid2d(8192, 8192, seed=0x5EED0004)
        info = G.info
        dist, st = ctx.sssp(G, 0, AT=G)
        d = dist.download(np.float32, info.n_rows)
        begins, cnt = pc.sample_blocks(info.n_rows, 1 << 14)
        ok = bool(st.converged) and all(pc.sssp_rows_ok(d, *G.download(b, cnt), b, 0) for b in begins)
        if not ok:
            raise SystemExit("bench: SSSP distances are not the Bellman fixed point -- refusing to report a number")
        r = {"graph": "2-D grid 8192x8192 4-neighbour, integer weights 1..255", "n": info.n_rows, "nnz": info.nnz,
             "rounds": st.iterations, "converged": bool(st.converged), "ms": st.total_ns / 1e6,
             "gteps": st.edges_traversed / st.total_ns, "parity_checked": True, "dist_checksum64": pc.checksum64(d),
             "parity": f"Bellman fixed-point condition in float arithmetic on {len(begins) * cnt} sampled rows"}
        del d
        # one dense min-plus sweep over the same operator (the reference's per-iteration launch, app/sssp.cpp:112-153):
        # gathers are local here, so this is the HBM-bound face of the same SpMV kernels (SURVEY.md 8d, C4 bytes/sweep)
        out = ctx.empty(info.n_rows * 4)
        ts = [ctx.spmv(G, shb.MIN_PLUS, dist, dist, 0.0, 0.0, out) for _ in range(12)][2:]
        ms = float(np.median(ts)) / 1e6
        peak, _ = measured_peak_gbs()
        r["dense_sweep"] = {"ms": ms, "algorithmic_bytes": info.algorithmic_bytes_spmv,
                            "effective_gbs": info.algorithmic_bytes_spmv / ms / 1e6,
                            "frac_of_measured_hbm_peak": info.algorithmic_bytes_spmv / ms / 1e6 / peak,
                            "kernel": f"variant={info.auto_variant} lanes={info.auto_lanes} (auto), min-plus, y = x"}
        out.free()
        dist.free()
        G.free()
        return r

    def pagerank():
        A = ctx.synth_rmat(24, 16, seed=0x5EED0026, symmetrise=False, value_mode=1)
        info = A.info
        res, prev, st = ctx.pagerank(A, delta=0.0, max_iters=20, check_every=20, both=True)  # fixed 20 sweeps (8d C5)
        xn, xp = res.download(np.float32, info.n_rows), prev.download(np.float32, info.n_rows)
        res.free(); prev.free()
        beta = float((np.float32(1.0) - np.float32(0.85)) / np.float32(info.n_rows))
        begins, cnt = pc.sample_blocks(info.n_rows, 4096)
        err = max(pc.pagerank_rows_err(orc, xp, xn[b:b + cnt], *A.download(b, cnt), beta) for b in begins)
        if not err <= 1e-5:
            raise SystemExit(f"bench: PageRank sweep differs from the oracle (rel {err:.3g}) -- refusing to report a number")
        r = {"parity_checked": True, "parity_rel_err": err,"graph": "RMAT scale 24 edgefactor 16 directed (config 5 family; scale 26 needs the multi-GPU partition)",
             "n": info.n_rows, "nnz": info.nnz, "sweeps": st.iterations, "ms": st.total_ns / 1e6,
             "gteps": info.nnz * st.iterations / st.total_ns,
             "effective_gbs": info.algorithmic_bytes_spmv * st.iterations / st.total_ns}
        A.free()
        return r

    guarded("bfs", bfs)
    guarded("sssp", sssp)
    guarded("pagerank", pagerank)
    return out


def partitioned_extras(ctx, rank, world):
    """GTEPS of the row-partitioned drivers at this GPU count: PageRank on config 5 (RMAT scale 26,
    2^30 non-zeros, 20 sweeps) and, for N > 1, BFS on config 3 (RMAT scale 24) and SSSP on config 4 (8192^2 grid).
    The work runs through the boundary's own multi-GPU drivers, shb_group_* (include/shb.h): ONE process -- rank 0 --
    drives all N devices with peer access; the other torchrun ranks wait on the rendezvous store (a CPU wait: no NCCL
    kernel may sit on a device while rank 0's cooperative kernels need all of its SMs).  Should the group path fail,
    every rank falls back to the torch.distributed drivers (tools/multi_*.py) and the line says so.  Every number is
    gated by a parity check; a mismatch aborts the run.  Every rank must call this."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    out = {}
    store = None
    if world > 1:
        import torch.distributed as dist
        store = dist.distributed_c10d._get_default_store()
    verdict = None
    if rank == 0:
        try:
            import group_bench
            import sparseharness_b200 as shb
            ctx.synchronize()
            g = shb.Group(world)
            res = {"pagerank_config5": group_bench.pagerank(g)}
            if world > 1:
                res["bfs_partitioned"] = group_bench.bfs(g)
                res["sssp_partitioned"] = group_bench.sssp(g)
            g.close()
            verdict = {"ok": True, "res": res}
        except Exception as e:  # fall back to the torch.distributed drivers below
            verdict = {"ok": False, "error": repr(e)[:400]}
        if store is not None:
            store.set("shb_group_verdict", json.dumps(verdict))
    else:
        verdict = json.loads(store.get("shb_group_verdict").decode())  # blocks on the CPU until rank 0 is done
    if verdict["ok"]:
        for name, r in verdict["res"].items():
            out[name] = r
            if not r.get("parity_ok", False):
                raise SystemExit(f"bench: {name} (shb_group) failed its parity gate at {world} GPUs -- refusing to report a number")
        return out
    out["group_path_error"] = verdict.get("error")

    def gate(name, r):
        r["driver"] = "torch.distributed fallback (tools/multi_*.py)"
        out[name] = r
        if not r.get("parity_ok", False):  # identical on every rank (all-reduced): all ranks leave together
            raise SystemExit(f"bench: {name} failed its parity gate at {world} GPUs -- refusing to report a number")

    try:
        import multi_pagerank
        r, _, _ = multi_pagerank.run(26, 16, 20, "fused", 0.0, "sampled", repeat=2, ctx=ctx)
        gate("pagerank_config5", r)
    except SystemExit:
        raise
    except Exception as e:
        out["pagerank_config5"] = {"error": repr(e)[:300]}
    if world > 1:
        try:
            import multi_bfs
            r, _, _ = multi_bfs.run(24, 16, "sampled", repeat=3, ctx=ctx)
            gate("bfs_partitioned", r)
        except SystemExit:
            raise
        except Exception as e:
            out["bfs_partitioned"] = {"error": repr(e)[:300]}
        try:
            import multi_sssp
            r, _, _ = multi_sssp.run(8192, 8192, "sampled", repeat=2, ctx=ctx)
            gate("sssp_partitioned", r)
        except SystemExit:
            raise
        except Exception as e:
            out["sssp_partitioned"] = {"error": repr(e)[:300]}
    return out


def run_ours(args):
    # stdout carries the ONE JSON line and nothing else: libraries that write to fd 1 (NCCL prints its version
    # there) are pointed at stderr for the whole run; the line itself goes to the saved descriptor
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        torch.cuda.set_device(local)
    n_gpus = world

    import sparseharness_b200 as shb  # raises if libshb.so is missing: no fallback path exists
    from oracle import oracle as orc  # checker + cpu_baseline only

    ctx = shb.Context(local)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- inputs resident in HBM
    n_rows_total = ROWS_PER_GPU * world
    A = ctx.synth_uniform(n_rows_total, N_COLS, NNZ_PER_ROW, seed=0x5EED0002, val_mode=0,
                          row_begin=rank * ROWS_PER_GPU, row_count=ROWS_PER_GPU)
    info = A.info
    assert info.algorithmic_bytes_spmv == ALGO_BYTES, (info.algorithmic_bytes_spmv, ALGO_BYTES)
    x = ctx.synth_vector(N_COLS, seed=0x5EED0003, mode=0)
    out = ctx.empty(ROWS_PER_GPU * 4)
    cfg = shb.launch_cfg()  # SHB_VARIANT_AUTO
    W = max(args.warmup, 3)
    K = args.steps

    # ---- device-resident throughput ("value")
    for _ in range(W):
        ctx.spmv(A, shb.PLUS_TIMES, x, None, 1.0, 0.0, out, cfg, timed=False)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        ev0.record()
        for _ in range(K):
            ctx.spmv(A, shb.PLUS_TIMES, x, None, 1.0, 0.0, out, cfg, timed=False)
        ev1.record()
        barrier()
        dev_ms = ev0.elapsed_time(ev1)

        # ---- end to end through the host-vector C-ABI call: x from pinned host memory, out back to it
        xh = torch.empty(N_COLS, dtype=torch.float32, pin_memory=True)
        oh = torch.empty(ROWS_PER_GPU, dtype=torch.float32, pin_memory=True)
        xh.copy_(torch.from_numpy(x.download(np.float32, N_COLS)))
        xh_np, oh_np = xh.numpy(), oh.numpy()
        Ke = max(6, min(K, 50))
        for _ in range(3):
            ctx.spmv_host(A, shb.PLUS_TIMES, xh_np, None, 1.0, 0.0, oh_np, cfg)
        barrier()
        t0 = time.perf_counter()
        e2e_dev_ns = 0
        for _ in range(Ke):
            tot, _ker = ctx.spmv_host(A, shb.PLUS_TIMES, xh_np, None, 1.0, 0.0, oh_np, cfg)
            e2e_dev_ns += tot
        barrier()
        e2e_single_s = time.perf_counter() - t0
        # the same Ke steps as ONE pipelined batch call: step i+1's upload and step i-1's download overlap SpMV i
        xh2 = torch.empty(N_COLS, dtype=torch.float32, pin_memory=True)
        oh2 = torch.empty(ROWS_PER_GPU, dtype=torch.float32, pin_memory=True)
        xh2.copy_(xh)
        xs_b = [xh_np if i % 2 == 0 else xh2.numpy() for i in range(Ke)]
        os_b = [oh_np if i % 2 == 0 else oh2.numpy() for i in range(Ke)]
        ctx.spmv_host_batch(A, shb.PLUS_TIMES, xs_b[:4], None, 1.0, 0.0, os_b[:4], cfg)
        barrier()
        t0 = time.perf_counter()
        ctx.spmv_host_batch(A, shb.PLUS_TIMES, xs_b, None, 1.0, 0.0, os_b, cfg)
        barrier()
        e2e_s = time.perf_counter() - t0
        batch_ok = bool(np.array_equal(oh2.numpy()[:4096], oh_np[:4096]))
    times = torch.tensor([dev_ms, e2e_s * 1e3, e2e_single_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, e2e_single_ms = [float(v) for v in times.tolist()]

    # ---- parity spot check outside the timed region (the oracle is the checker)
    got = out.download(np.float32, 4096)
    rp, col, val = orc.gen_uniform(n_rows_total, N_COLS, NNZ_PER_ROW, row_begin=rank * ROWS_PER_GPU, row_count=4096)
    xs = orc.gen_vector(N_COLS)
    want = orc.spmv(orc.PLUS_TIMES, rp, col, val, xs, np.zeros(4096, np.float32), 1.0, 0.0)
    parity_ok = bool(np.abs(got - want).max() <= 1e-5 * np.abs(want).max()) and bool(np.array_equal(oh_np[:4096], got)) and batch_ok
    if not parity_ok:
        raise SystemExit("bench: SpMV result differs from the oracle -- refusing to report a number")

    ms_per_step = dev_ms / K
    value = ALGO_BYTES * n_gpus / (ms_per_step * 1e-3) / 1e9
    per_gpu = ALGO_BYTES / (ms_per_step * 1e-3) / 1e9
    peak, peak_src = measured_peak_gbs()
    e2e_value = ALGO_BYTES * n_gpus / (e2e_ms / Ke * 1e-3) / 1e9
    clocks = clk.summary()

    # the kernel the AUTO schedule resolved to on this operator (never a literal: the line must name what ran)
    fam = {1: "spmv_scalar_kernel", 2: "spmv_vector_kernel", 3: "spmv_stream_kernel", 4: "pb_fused_kernel"}.get(info.auto_variant, "?")
    kernel_name = {1: f"{fam}<PlusTimesF32,true>", 2: f"{fam}<PlusTimesF32,{info.auto_lanes},2,true,true>",
                   3: f"{fam}<PlusTimesF32,true,true> + spmv_fixup_kernel", 4: f"{fam}<PlusTimesF32,true>"}.get(info.auto_variant, fam)
    binding = ("not HBM on this operator: every one of the 268M uniform-random x gathers is its own L1TEX->XBAR request, one per clock "
               "per SM -- a kernel doing ONLY those gathers takes 0.937 ms on this GPU (profiles/r02_gather_microbench.jsonl; ncu on this "
               "kernel: l1tex__m_l1tex2xbar_req_cycles_active 90%, lts__throughput 85%, dram 42%, profiles/r02_summary.md); the same "
               "kernels with local gathers are HBM-bound, see roofline_local_gathers") if info.auto_variant != 4 else \
              ("propagation blocking: x slices gathered from shared memory, products through an L2-resident ring, every global "
               "access a coalesced 128-bit stream (sparseharness_b200/csrc/pb.cuh; profiles/r02_summary.md)")
    line = {
        "metric": "spmv_effective_hbm_gbs", "value": value, "unit": "GB/s", "n_gpus": n_gpus, "steps": K, "warmup": W,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "rows_per_gpu": ROWS_PER_GPU, "nnz_per_gpu": ROWS_PER_GPU * NNZ_PER_ROW,
                   "algorithmic_bytes_per_step_per_gpu": ALGO_BYTES, "parallelism": f"rows x{n_gpus}, x replicated, no collective",
                   "kernel": f"variant={info.auto_variant} lanes={info.auto_lanes} (SHB_VARIANT_AUTO)",
                   "l2": "inputs 2.35 GB per GPU >> 126 MB L2, no flush between steps (x, 64 MB, may stay L2 resident)"},
        "gflops": FLOPS * n_gpus / (ms_per_step * 1e-3) / 1e9,
        "roofline": {"bound": "hbm", "achieved": per_gpu, "peak": peak, "unit": "GB/s", "frac": per_gpu / peak,
                     "traffic": None, "peak_source": peak_src, "kernel": kernel_name,
                     "note": "achieved = 2348810244 B / mean launch time over the timed region (CUDA events on the launch stream)",
                     "binding_unit": binding},
        "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": N_COLS * 4, "d2h_bytes_per_step": ROWS_PER_GPU * 4,
                "ms_per_step": e2e_ms / Ke, "steps": Ke,
                "api": "shb_spmv_host_batch: one call for all steps; every step uploads its x from pinned host memory and "
                       "downloads its result to pinned host memory; upload i+1 / SpMV i / download i-1 overlap on three "
                       "streams; operator resident; host wall clock, max over ranks",
                "single_call": {"value": ALGO_BYTES * n_gpus / (e2e_single_ms / Ke * 1e-3) / 1e9, "unit": "GB/s",
                                "ms_per_step": e2e_single_ms / Ke,
                                "api": "shb_spmv_host, one blocking call per step: upload, row slabs whose results go back "
                                       "on a second stream while the next slab computes"}},
        "gpu_launches": K * n_gpus, "clocks": clocks, "parity_checked": parity_ok,
    }
    if info.auto_variant != 4:
        # the two units that do bind this operator, from the same measured launch time (context for `frac`, not a claim):
        # one L1TEX->XBAR request per gather (its own 128-byte line) + per streamed line, one per clock per SM;
        # one 32-byte L2 sector per gather + the streamed bytes, ~6300 B per clock chip-wide (microarchitecture guide)
        try:
            import torch as _t
            n_sm = _t.cuda.get_device_properties(_t.cuda.current_device()).multi_processor_count
        except Exception:  # noqa: BLE001
            n_sm = 148
        hz = float(clocks.get("sm_mhz") or 1965.0) * 1e6 if isinstance(clocks, dict) else 1965e6
        nnz_g = ROWS_PER_GPU * NNZ_PER_ROW
        stream_bytes = ALGO_BYTES - N_COLS * 4
        t = ms_per_step * 1e-3
        req = nnz_g + stream_bytes / 128.0
        sec = nnz_g + stream_bytes / 32.0
        line["roofline"]["request_ceilings"] = {
            "l1tex_to_xbar_requests": {"achieved_per_s": req / t, "peak_per_s": n_sm * hz, "frac": req / t / (n_sm * hz)},
            "l2_sectors": {"achieved_per_s": sec / t, "peak_per_s": 6300.0 / 32.0 * hz, "frac": sec / t / (6300.0 / 32.0 * hz)},
            "pure_gather_ms": 0.937,
            "note": "derived from ms_per_step and the sampled SM clock; profiles/r02_summary.md section 1 has the ncu counters. "
                    "pure_gather_ms: measured time of a kernel doing ONLY this operator's 268M random gathers from a 64 MB table "
                    "(tools/microbench/gather_bench.cu, profiles/r02_gather_microbench.jsonl): the floor of any row-major schedule"}
    traffic_file = os.path.join(ROOT, "profiles", "traffic_bytes_per_launch.json")
    if os.path.exists(traffic_file):
        try:
            tf = json.load(open(traffic_file))
            # only the counters of THIS kernel family may be attached (tools/ncu_traffic.py writes name + git head)
            if tf.get("kernel", "").split("<")[0].split("(")[0].strip().endswith(fam):
                line["roofline"]["traffic"] = tf["dram_bytes_per_launch"]
                line["roofline"]["traffic_source"] = {"kernel": tf.get("kernel"), "git_head": tf.get("git_head"), "file": "profiles/traffic_bytes_per_launch.json"}
            else:
                line["roofline"]["traffic_note"] = f"no ncu capture of {fam} on file (profiles/traffic_bytes_per_launch.json holds {tf.get('kernel')})"
        except Exception:
            pass

    if rank == 0 and n_gpus == 1:
        rp, col, val, xs, nbytes = cpu_sample_inputs(orc)
        threads = orc.num_threads()
        t = time_cpu(orc, rp, col, val, xs, threads, steps=max(3, min(40, args.cpu_steps)), warmup=1)
        best = float(np.min(t))
        line["cpu_baseline"] = {"value": nbytes / best / 1e9, "unit": "GB/s", "cores": threads, "kind": "port",
                                "gflops": 2 * len(col) / best / 1e9, "ms_per_pass": best * 1e3,
                                "sample": f"first {CPU_SAMPLE_ROWS} rows ({len(col)} non-zeros) of the same operator, x full "
                                          f"length; oracle restatement of inc/spmv_gold.h:17-26, OpenMP over rows, best of {len(t)}"}
        t1 = time_cpu(orc, rp, col, val, xs, 1, steps=3, warmup=0)
        line["cpu_baseline"]["single_thread_gbs"] = nbytes / float(np.min(t1)) / 1e9
        del rp, col, val
        ref = time_reference_gold(orc, 1, 20)  # the unmodified reference gold on its bounded sample (~10 s)
        if ref is not None:
            port = line["cpu_baseline"]
            line["cpu_baseline"] = {"value": ref["bytes"] / ref["mean_s"] / 1e9, "unit": "GB/s", "cores": 1, "kind": "reference",
                                    "gflops": 2 * ref["nnz"] / ref["mean_s"] / 1e9, "ms_per_pass": ref["mean_s"] * 1e3,
                                    "sample": ref["sample"], "port_all_cores": port}
    A.free(); x.free(); out.free()
    if rank == 0 and n_gpus == 1 and args.extras:
        line["graph_gteps"] = graph_extras(ctx, shb, orc)
    if args.extras:
        # the iterative drivers on the row-partitioned operator (every rank takes part; rank 0 reports)
        part = partitioned_extras(ctx, rank, n_gpus)
        if rank == 0:
            line.setdefault("graph_gteps", {}).update(part)
    if rank == 0:
        gg = line.get("graph_gteps", {})
        pick = lambda *ks: next((gg[k].get("gteps") for k in ks if isinstance(gg.get(k), dict) and "gteps" in gg[k]), None)
        # the numbers that say something about the multi-GPU design (the headline SpMV shards with no collective)
        line["scaling_curve"] = {"n_gpus": n_gpus, "pagerank_config5_gteps": pick("pagerank_config5"),
                                 "bfs_config3_gteps": pick("bfs_partitioned", "bfs"),
                                 "sssp_config4_gteps": pick("sssp_partitioned", "sssp"),
                                 "parity_checked": all(bool(v.get("parity_checked")) for v in gg.values()
                                                       if isinstance(v, dict) and "gteps" in v)}
        ds = line.get("graph_gteps", {}).get("sssp", {}).get("dense_sweep")
        local_traffic = None
        try:
            local_traffic = json.load(open(traffic_file))["local_gathers"]["dram_bytes_per_launch"]
        except Exception:
            pass
        if ds:  # the HBM-bound face of the same kernels, as a second roofline object next to the headline one
            line["roofline_local_gathers"] = {
                "bound": "hbm", "achieved": ds["effective_gbs"], "peak": peak, "unit": "GB/s",
                "frac": ds["frac_of_measured_hbm_peak"], "traffic": local_traffic,
                "workload": "configs[3] operator (8192x8192 grid, 268M non-zeros), one dense min-plus sweep, "
                            f"{ds['algorithmic_bytes']} algorithmic bytes", "kernel": "spmv_stream_kernel<MinPlusF32> + spmv_fixup_kernel"}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=int(os.environ.get("WORLD_SIZE", "1")))
    ap.add_argument("--steps", type=int, default=60)  # (200 back-to-back launches run into the software power cap: 1.12 ms per step against 1.07)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--extras", type=int, default=1, help="also report BFS/SSSP/PageRank GTEPS on 1 GPU")
    ap.add_argument("--cpu-steps", type=int, default=12)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
