"""ctypes/numpy binding of oracle/_build/liboracle.so (the CPU oracle, shb_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Nothing under sparseharness_b200/ imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "liboracle.so")
REF_DRIVER = os.path.join(HERE, "_ref", "ref_driver")

PLUS_TIMES, MIN_PLUS, OR_AND, MAX_MIN = 0, 1, 2, 3
TERM_ABSDIFF, TERM_EQUAL = 0, 1
FLT_MAX = np.float32(3.4028235e38)
INT_MIN = np.int32(-2147483648)
INT_MAX = np.int32(2147483647)


def build(ref: bool = True) -> None:
    """Compile the oracle (and, when /root/reference is present, the reference driver)."""
    subprocess.check_call(["make", "-s", "-C", HERE, "all"])
    if ref:
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build(ref=False)
        _lib = C.CDLL(LIB_PATH)
        _lib.orc_zero_f32.restype = C.c_float
        _lib.orc_iterate_f32.restype = C.c_int64
        _lib.orc_iterate_i32.restype = C.c_int64
        _lib.orc_pagerank.restype = C.c_int64
        _lib.orc_gen_grid2d.restype = C.c_int64
        _lib.orc_splitmix64.restype = C.c_uint64
        _lib.orc_splitmix64.argtypes = [C.c_uint64]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    return None if a is None else np.ascontiguousarray(a, dtype=dt)


def write_mtx(path, n_rows, n_cols, row_ptr, col, val, second_index_is_row=False) -> None:
    """CSR block -> Matrix Market text file (what the reference's loader reads, src/sparse_matrix.cpp:11-70).
    second_index_is_row writes "col row value" lines: the reference takes the second index as the row (Q1), so
    that is the file from which ITS loader rebuilds exactly these rows."""
    rp, c, v = _c(row_ptr, np.int64), _c(col, np.int32), _c(val, np.float32)
    if lib().orc_write_mtx(path.encode(), C.c_int64(n_rows), C.c_int64(n_cols), _p(rp), _p(c), _p(v), C.c_int(int(second_index_is_row))) != 0:
        raise OSError(f"cannot write {path}")


REF_DRIVER = os.path.join(HERE, "_ref", "ref_driver")


def reference_gold_time(mtx_path, warmup, reps, timeout=600):
    """Times the UNMODIFIED reference's Gold<float>::spmv (inc/spmv_gold.h:9-28, single threaded, matrix passed by
    value) on a Matrix Market file through oracle/_ref/ref_driver.  Returns (mean_ns, best_ns) or None when the
    driver is not there (it is compiled from /root/reference by oracle/Makefile and travels prebuilt)."""
    if not os.path.exists(REF_DRIVER):
        return None
    r = subprocess.run([REF_DRIVER, "time", mtx_path, str(int(reps)), str(int(warmup))], stdout=subprocess.PIPE,
                       stderr=subprocess.DEVNULL, timeout=timeout)
    for line in reversed(r.stdout.decode(errors="replace").splitlines()):
        if line.startswith("TIME "):
            f = line.split()
            return float(f[f.index("mean_ns") + 1]), float(f[f.index("best_ns") + 1])
    return None


def num_threads() -> int:
    return int(lib().orc_num_threads())


class Coo(C.Structure):
    _fields_ = [("rows", C.c_int32), ("cols", C.c_int32), ("file_nnz", C.c_int64), ("n", C.c_int64),
                ("I", C.POINTER(C.c_int32)), ("J", C.POINTER(C.c_int32)), ("V", C.POINTER(C.c_double)),
                ("is_pattern", C.c_int), ("is_symmetric", C.c_int)]


def mm_load_csr(path: str, elem: str = "float", normalise: str = "none", truncate: bool = True):
    """Reference-faithful ingest: (n_rows, n_cols, row_ptr[int64], col[int32], val[float32|int32])."""
    et = 0 if elem == "float" else 1
    coo = Coo()
    rc = lib().orc_mm_read(path.encode(), et, C.byref(coo))
    if rc != 0:
        raise RuntimeError(f"orc_mm_read({path}) failed: {rc}")
    try:
        if normalise == "pr":
            lib().orc_pagerank_normalise_f32(C.byref(coo), C.c_float(0.85), C.c_float(0.0))
        elif normalise == "scc":
            lib().orc_scc_normalise_i32(C.byref(coo))
        row_ptr = np.zeros(coo.rows + 1, dtype=np.int64)
        col = np.zeros(coo.n, dtype=np.int32)
        val = np.zeros(coo.n, dtype=np.float32 if et == 0 else np.int32)
        rc = lib().orc_coo_to_csr(C.byref(coo), et, int(truncate), _p(row_ptr), _p(col), _p(val))
        if rc != 0:
            raise RuntimeError("orc_coo_to_csr failed")
        return int(coo.rows), int(coo.cols), row_ptr, col, val
    finally:
        lib().orc_coo_free(C.byref(coo))


def spmv(semiring, row_ptr, col, val, x, y, alpha, beta, order=0, threads=1, vlength=None):
    n = len(row_ptr) - 1
    row_ptr = _c(row_ptr, np.int64)
    col = _c(col, np.int32)
    vlength = len(x) if vlength is None else vlength
    if semiring in (PLUS_TIMES, MIN_PLUS):
        val = _c(val, np.float32); x = _c(x, np.float32); y = _c(y, np.float32)
        out = np.empty(n, dtype=np.float32)
        lib().orc_spmv_f32(semiring, C.c_int64(n), C.c_int64(vlength), _p(row_ptr), _p(col), _p(val), _p(x), _p(y),
                           C.c_float(alpha), C.c_float(beta), _p(out), order, threads)
    else:
        val = _c(val, np.int32); x = _c(x, np.int32); y = _c(y, np.int32)
        out = np.empty(n, dtype=np.int32)
        lib().orc_spmv_i32(semiring, C.c_int64(n), C.c_int64(vlength), _p(row_ptr), _p(col), _p(val), _p(x), _p(y),
                           C.c_int32(int(alpha)), C.c_int32(int(beta)), _p(out), order, threads)
    return out


def gold_spmv(row_ptr, col, val, x, y, alpha=1.0, beta=0.0, zero=0.0):
    n = len(row_ptr) - 1
    row_ptr = _c(row_ptr, np.int64); col = _c(col, np.int32); val = _c(val, np.float32)
    x = _c(x, np.float32); y = _c(y, np.float32)
    out = np.empty(n, dtype=np.float32)
    lib().orc_gold_spmv_f32(C.c_int64(n), _p(row_ptr), _p(col), _p(val), _p(x), _p(y), C.c_int64(len(y)),
                            C.c_float(alpha), C.c_float(beta), C.c_float(zero), _p(out))
    return out


def iterate(semiring, row_ptr, col, val, x0, y0, alpha, beta, delta=1e-4, term=None, y_rebind=True,
            max_iters=1 << 30, order=0, threads=1):
    """Faithful fixed-point driver (app/*.cpp executeRun).  Returns (result, n_kernel_launches)."""
    n = len(row_ptr) - 1
    row_ptr = _c(row_ptr, np.int64); col = _c(col, np.int32)
    if semiring in (PLUS_TIMES, MIN_PLUS):
        term = TERM_ABSDIFF if term is None else term
        val = _c(val, np.float32); x0 = _c(x0, np.float32); y0 = _c(y0, np.float32)
        res = np.empty(n, dtype=np.float32)
        it = lib().orc_iterate_f32(semiring, C.c_int64(n), _p(row_ptr), _p(col), _p(val), _p(x0), _p(y0),
                                   C.c_float(alpha), C.c_float(beta), C.c_double(delta), term, int(y_rebind),
                                   C.c_int64(max_iters), order, threads, _p(res))
    else:
        term = TERM_EQUAL if term is None else term
        val = _c(val, np.int32); x0 = _c(x0, np.int32); y0 = _c(y0, np.int32)
        res = np.empty(n, dtype=np.int32)
        it = lib().orc_iterate_i32(semiring, C.c_int64(n), _p(row_ptr), _p(col), _p(val), _p(x0), _p(y0),
                                   C.c_int32(int(alpha)), C.c_int32(int(beta)), term, int(y_rebind),
                                   C.c_int64(max_iters), order, threads, _p(res))
    return res, int(it)


def bfs_levels(row_ptr, col, val=None, root=0):
    n = len(row_ptr) - 1
    row_ptr = _c(row_ptr, np.int64); col = _c(col, np.int32); val = _c(val, np.int32)
    level = np.empty(n, dtype=np.int32)
    edges = C.c_int64(0)
    nlev = lib().orc_bfs_levels(C.c_int64(n), _p(row_ptr), _p(col), _p(val), C.c_int64(root), _p(level), C.byref(edges))
    return level, int(nlev), int(edges.value)


def sssp(row_ptr, col, val, source=0):
    n = len(row_ptr) - 1
    row_ptr = _c(row_ptr, np.int64); col = _c(col, np.int32); val = _c(val, np.float32)
    dist = np.empty(n, dtype=np.float32)
    lib().orc_sssp_dijkstra(C.c_int64(n), _p(row_ptr), _p(col), _p(val), C.c_int64(source), _p(dist))
    return dist


def pagerank(row_ptr, col, val, damping=0.85, delta=1e-4, max_iters=1000, acc_double=False, threads=1):
    n = len(row_ptr) - 1
    row_ptr = _c(row_ptr, np.int64); col = _c(col, np.int32); val = _c(val, np.float32)
    res = np.empty(n, dtype=np.float32)
    x_init = np.float32(1.0) / np.float32(n)
    beta = (np.float32(1.0) - np.float32(damping)) / np.float32(n)
    it = lib().orc_pagerank(C.c_int64(n), _p(row_ptr), _p(col), _p(val), C.c_float(x_init), C.c_float(beta),
                            C.c_double(delta), C.c_int64(max_iters), int(acc_double), threads, _p(res))
    return res, int(it)


def power_iteration(row_ptr, col, val, x0, delta=1e-4, max_iters=10, order=0):
    """Canonical eigenvector driver: (x_k, eigenvalue estimate, iterations)."""
    n = len(row_ptr) - 1
    row_ptr = _c(row_ptr, np.int64); col = _c(col, np.int32); val = _c(val, np.float32); x0 = _c(x0, np.float32)
    res = np.empty(n, dtype=np.float32)
    ev = C.c_float(0)
    f = lib().orc_power_iteration
    f.restype = C.c_int64
    it = f(C.c_int64(n), _p(row_ptr), _p(col), _p(val), _p(x0), C.c_double(delta), C.c_int64(max_iters), order, _p(res), C.byref(ev))
    return res, float(ev.value), int(it)


def pagerank_values(row_ptr, col, damping=0.85):
    n = len(row_ptr) - 1
    row_ptr = _c(row_ptr, np.int64); col = _c(col, np.int32)
    val = np.empty(int(row_ptr[-1]), dtype=np.float32)
    lib().orc_pagerank_values(C.c_int64(n), _p(row_ptr), _p(col), C.c_float(damping), _p(val))
    return val


# ---- synthetic workloads -------------------------------------------------
SEED_C2, SEED_C2_X, SEED_C3, SEED_C4, SEED_C5 = 0x5EED0002, 0x5EED0003, 0x5EED0024, 0x5EED0004, 0x5EED0026


def gen_uniform(n_rows, n_cols, k=16, seed=SEED_C2, val_mode=0, row_begin=0, row_count=None):
    row_count = n_rows - row_begin if row_count is None else row_count
    row_ptr = np.empty(row_count + 1, dtype=np.int64)
    col = np.empty(row_count * k, dtype=np.int32)
    val = np.empty(row_count * k, dtype=np.float32)
    lib().orc_gen_uniform(C.c_int64(n_rows), C.c_int64(n_cols), k, C.c_uint64(seed), val_mode,
                          C.c_int64(row_begin), C.c_int64(row_count), _p(row_ptr), _p(col), _p(val))
    return row_ptr, col, val


def gen_vector(n, seed=SEED_C2_X, mode=0):
    x = np.empty(n, dtype=np.float32)
    lib().orc_gen_vector(C.c_int64(n), C.c_uint64(seed), mode, _p(x))
    return x


def gen_rmat(scale, edgefactor=16, seed=SEED_C3, symmetrise=True):
    n = 1 << scale
    n_edges = n * edgefactor
    row_ptr = np.empty(n + 1, dtype=np.int64)
    nnz = C.c_int64(0)
    cap = 2 * n_edges if symmetrise else n_edges
    col = np.empty(cap, dtype=np.int32)
    lib().orc_gen_rmat_csr(scale, C.c_int64(n_edges), C.c_uint64(seed), int(symmetrise), _p(row_ptr), _p(col),
                           C.byref(nnz))
    return row_ptr, col[: nnz.value].copy()


def gen_grid2d(w, h, seed=SEED_C4):
    n = w * h
    row_ptr = np.empty(n + 1, dtype=np.int64)
    nnz = 4 * n - 2 * w - 2 * h
    col = np.empty(nnz, dtype=np.int32)
    val = np.empty(nnz, dtype=np.float32)
    got = lib().orc_gen_grid2d(C.c_int64(w), C.c_int64(h), C.c_uint64(seed), _p(row_ptr), _p(col), _p(val))
    assert got == nnz, (got, nnz)
    return row_ptr, col, val


# ---- the real reference (oracle/_ref/ref_driver), when built ------------------------------
def ref_available() -> bool:
    return os.path.exists(REF_DRIVER)


def _run_ref(*args) -> str:
    return subprocess.run([REF_DRIVER, *map(str, args)], check=True, stdout=subprocess.PIPE,
                          stderr=subprocess.DEVNULL, text=True).stdout


def ref_rows(mtx, elem="float", normalise="none"):
    """Loader's ragged rows (after Q1-Q3) as CSR arrays, straight from the reference binary."""
    lines = _run_ref("rows", mtx, elem, normalise).splitlines()
    i = next(k for k, l in enumerate(lines) if l.startswith("ROWS "))
    _, h, w, nz, nrows = lines[i].split()
    rp = [0]; cols = []; vals = []
    for l in lines[i + 1: i + 1 + int(nrows)]:
        t = l.split()
        n = int(t[0])
        cols.extend(int(c) for c in t[1:1 + 2 * n:2])
        vals.extend(float(v) for v in t[2:2 + 2 * n:2])
        rp.append(rp[-1] + n)
    vd = np.float32 if elem == "float" else np.int32
    return int(h), int(w), np.array(rp, np.int64), np.array(cols, np.int32), np.array(vals, np.float64).astype(vd)


def ref_gold(mtx, xmode=0):
    lines = _run_ref("gold", mtx, xmode).splitlines()
    i = next(k for k, l in enumerate(lines) if l.startswith("GOLD "))
    n = int(lines[i].split()[1])
    return np.array([float(v) for v in lines[i + 1: i + 1 + n]], dtype=np.float64).astype(np.float32)


def ref_time_gold(mtx, reps=5):
    out = _run_ref("time", mtx, reps)
    t = next(l for l in out.splitlines() if l.startswith("TIME ")).split()
    return {"best_ns": float(t[2]), "mean_ns": float(t[4]), "reps": int(t[6])}


REF_KERNELS = os.path.join(HERE, "_ref", "ref_kernels")
REF_KERNEL_ALGS = ("spmv", "sssp", "bfs", "pr", "scc")


def ref_kernel_run(mtx, alg, variant, cap=40, delta=1e-4, gsz=1, xmode=0):
    """The reference's OWN kernel text (example/<alg>/kernel{5,6}.json:3) run on the CPU over the reference's OWN
    cl_encode blobs with the app's initial state: one launch (`step1`) and the executeRun recurrence capped at
    `cap` launches (oracle/ref_kernels.cpp).  Returns dict(h, w, iters, terminated, step1, final)."""
    import tempfile
    dt = np.float32 if alg in ("spmv", "sssp", "pr") else np.int32
    with tempfile.NamedTemporaryFile("r", suffix=".txt") as f:
        subprocess.run([REF_KERNELS, str(mtx), alg, str(variant), str(cap), repr(float(delta)), str(gsz), str(xmode),
                        f.name], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        lines = f.read().split()
    assert lines[0] == "H" and lines[2] == "W" and lines[4] == "ITERS" and lines[6] == "TERMINATED"
    out = {"h": int(lines[1]), "w": int(lines[3]), "iters": int(lines[5]), "terminated": int(lines[7])}
    i = 8
    for tag in ("STEP1", "FINAL"):
        assert lines[i] == tag
        n = int(lines[i + 1])
        words = np.array([int(t, 16) for t in lines[i + 2: i + 2 + n]], dtype=np.uint32)
        out[tag.lower()] = words.view(dt)
        i += 2 + n
    return out


def xmode_vector(n, xmode):
    i = np.arange(n, dtype=np.int64)
    return np.ones(n, np.float32) if xmode == 0 else (((7 * i + 3) % 11) - 5).astype(np.float32)
