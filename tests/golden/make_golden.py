"""Generate tests/golden/* from the UNMODIFIED reference (oracle/_ref/ref_driver).

Run in the build container (needs /root/reference):   python tests/golden/make_golden.py
Writes, per fixture matrix:
  tests/golden/mtx/<suitesparse-name>.mtx.gz   the public SuiteSparse input the reference ships as
                                               example/matrix{,2,3,4,5}.mtx (data, not code)
  tests/golden/ref_<name>.npz                  what the reference's own code produced for it:
      f_row_ptr/f_col/f_val   SparseMatrix<float> loader rows  (src/sparse_matrix.cpp:11-119)
      i_row_ptr/i_col/i_val   SparseMatrix<int>   loader rows
      scc_val                 values after scc_normalise      (src/sparse_matrix.cpp:433-456)
      pr_val                  values after pagerank_normalise (src/sparse_matrix.cpp:409-431; all
                              truncated to 0 by Q2 -- that IS the reference's behaviour)
      gold_x0 / gold_x1       Gold<float>::spmv (inc/spmv_gold.h:9-28), x == 1 and x[i]=((7i+3)%11)-5
      k{5,6}_<alg>_step1      ONE launch of the reference's own device kernel text, example/<alg>/kernel5.json:3
                              (glb-sdp: padded ELL, reversed reduction order) and kernel6.json:3 (glb-sdp-rsa:
                              ragged, forward order), executed on the CPU by oracle/_ref/ref_kernels over the blobs of
                              the reference's own cl_encode (src/sparse_matrix.cpp:122-399) with the app's initial
                              state; <alg> in spmv (x == 1), spmvx1 (x[i]=((7i+3)%11)-5), sssp, bfs, pr, scc
      k{5,6}_<alg>_final / _iters / _terminated
                              the executeRun recurrence (app/sssp.cpp:112-153 and twins) run to its own termination
                              test or to the cap in ITER_CAP (the reference has no cap; its bfs / pr loops need not stop)
"""
import gzip
import os
import shutil
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402

REF = "/root/reference/example"
FIXTURES = {  # reference file -> SuiteSparse name
    "matrix.mtx": "1138_bus",
    "matrix2.mtx": "ca-AstroPh",
    "matrix3.mtx": "Trefethen_20",
    "matrix4.mtx": "GD96_b",
    "matrix5.mtx": "arc130",
}


ITER_CAP = {"sssp": 400, "bfs": 40, "pr": 40, "scc": 40}
DELTA = 1e-4  # inc/common.h:27-29 default of -c/--delta


def kernel_goldens(src):
    """Arrays from the reference's own kernel sources (oracle/ref_kernels.cpp)."""
    out = {}
    for variant in (5, 6):
        for alg, xmode, tag in (("spmv", 0, "spmv"), ("spmv", 1, "spmvx1"), ("sssp", 0, "sssp"), ("bfs", 0, "bfs"),
                                ("pr", 0, "pr"), ("scc", 0, "scc")):
            cap = ITER_CAP.get(alg, 1)
            r = orc.ref_kernel_run(src, alg, variant, cap=cap, delta=DELTA, gsz=1, xmode=xmode)
            # an NDRange of 7 work-items gives the same bits as one (rows are independent)
            r7 = orc.ref_kernel_run(src, alg, variant, cap=cap, delta=DELTA, gsz=7, xmode=xmode)
            assert r7["iters"] == r["iters"] and r7["final"].tobytes() == r["final"].tobytes()
            out[f"k{variant}_{tag}_step1"] = r["step1"]
            if alg != "spmv":
                out[f"k{variant}_{tag}_final"] = r["final"]
                out[f"k{variant}_{tag}_iters"] = np.array([r["iters"], r["terminated"], cap], np.int64)
    return out


def main():
    orc.build(ref=True)
    out = os.path.join(ROOT, "tests", "golden")
    os.makedirs(os.path.join(out, "mtx"), exist_ok=True)
    for ref_name, name in FIXTURES.items():
        src = os.path.join(REF, ref_name)
        with open(src, "rb") as fi, gzip.GzipFile(os.path.join(out, "mtx", name + ".mtx.gz"), "wb", mtime=0) as fo:
            shutil.copyfileobj(fi, fo)
        h, w, frp, fc, fv = orc.ref_rows(src, "float", "none")
        _, _, irp, ic, iv = orc.ref_rows(src, "int", "none")
        _, _, srp, sc, sv = orc.ref_rows(src, "int", "scc")
        _, _, prp, pc, pv = orc.ref_rows(src, "float", "pr")
        assert (srp == irp).all() and (sc == ic).all() and (prp == frp).all() and (pc == fc).all()
        g0 = orc.ref_gold(src, 0)
        g1 = orc.ref_gold(src, 1)
        kg = kernel_goldens(src)
        np.savez_compressed(os.path.join(out, f"ref_{name}.npz"), shape=np.array([h, w], np.int64),
                            f_row_ptr=frp, f_col=fc, f_val=fv, i_row_ptr=irp, i_col=ic, i_val=iv,
                            scc_val=sv, pr_val=pv, gold_x0=g0, gold_x1=g1, **kg)
        print("   kernels:", {k: (int(v[0]), int(v[1])) for k, v in kg.items() if k.endswith("_iters")})
        print(f"{name}: {h}x{w} stored={len(fc)} maxrow={int(np.diff(frp).max())} gold0[:8]={g0[:8]} sum={g0.astype(np.float64).sum()}")


if __name__ == "__main__":
    main()
