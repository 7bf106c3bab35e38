"""CPU tests of the C++ harness host (host/): ingest + argument encoding + the in-run gold, through
just_parser_harness (no device needed), against the golden vectors produced by the UNMODIFIED reference
(tests/golden/ref_*.npz, see tests/golden/make_golden.py)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import FIXTURES, GOLDEN, ROOT

BIN = os.path.join(ROOT, "host", "bin")
KERNELS = os.path.join(GOLDEN, "kernels")
NAMES = list(FIXTURES.values())


def _have_host():
    return os.path.exists(os.path.join(BIN, "just_parser_harness"))


pytestmark = pytest.mark.skipif(not _have_host(), reason="host/bin not built (python __graft_entry__.py build)")


def _parse(tmp_path, mtx, kernel, elem="float", normalise="none", extra=()):
    prefix = str(tmp_path / "csr")
    env = dict(os.environ, SHB_QUIET_LOG="1")
    r = subprocess.run([os.path.join(BIN, "just_parser_harness"), "-i", "1", "-m", mtx, "-k", kernel, "--elem", elem,
                        "--normalise", normalise, "--dump_csr", prefix, *extra], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    vt = np.float32 if elem == "float" else np.int32
    info = dict(zip(*[iter(next(l for l in r.stdout.splitlines() if l.startswith("semiring ")).split())] * 2))
    return {"rp": np.fromfile(prefix + ".rowptr", np.int64), "col": np.fromfile(prefix + ".col", np.int32),
            "val": np.fromfile(prefix + ".val", vt), "x": np.fromfile(prefix + ".x", vt),
            "y": np.fromfile(prefix + ".y", vt), "gold": np.fromfile(prefix + ".gold", vt),
            "info": {k: int(v) for k, v in info.items()}, "stdout": r.stdout, "stderr": r.stderr}


@pytest.mark.parametrize("name", NAMES)
def test_host_loader_rows_and_gold_match_reference(tmp_path, mtx_dir, name):
    g = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))
    h = int(g["shape"][0])
    d = _parse(tmp_path, mtx_dir[name], os.path.join(KERNELS, "spmv", "glb-sdp-rsa.json"))
    assert len(d["rp"]) == h + 1  # no chunking -> no padded rows
    assert np.array_equal(d["rp"], g["f_row_ptr"]) and np.array_equal(d["col"], g["f_col"])
    assert np.array_equal(d["val"].view(np.int32), g["f_val"].view(np.int32))
    assert np.array_equal(d["gold"].view(np.int32), g["gold_x0"].view(np.int32))
    d = _parse(tmp_path, mtx_dir[name], os.path.join(KERNELS, "bfs", "glb-sdp.json"), elem="int")
    assert np.array_equal(d["rp"], g["i_row_ptr"]) and np.array_equal(d["col"], g["i_col"])
    assert np.array_equal(d["val"], g["i_val"])


@pytest.mark.parametrize("name", NAMES)
def test_host_normalisations_match_reference(tmp_path, mtx_dir, name):
    g = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))
    d = _parse(tmp_path, mtx_dir[name], os.path.join(KERNELS, "scc", "glb-sdp.json"), elem="int", normalise="scc")
    assert np.array_equal(d["val"], g["scc_val"])
    d = _parse(tmp_path, mtx_dir[name], os.path.join(KERNELS, "spmv", "glb-sdp.json"), normalise="pr")
    assert np.array_equal(d["val"].view(np.int32), g["pr_val"].view(np.int32))


def test_host_canonical_pagerank_values_match_oracle(tmp_path, mtx_dir, orc):
    # float values kept (canonical operator): same double-evaluated quotient as the oracle's restatement
    name = "GD96_b"
    d = _parse(tmp_path, mtx_dir[name], os.path.join(KERNELS, "spmv", "glb-sdp.json"), normalise="pr",
               extra=("--keep_values",))
    _, _, rp, col, val = orc.mm_load_csr(mtx_dir[name], "float", normalise="pr", truncate=False)
    assert np.array_equal(d["rp"], rp) and np.array_equal(d["col"], col)
    assert np.array_equal(d["val"].view(np.int32), val.view(np.int32))
    assert (d["val"] > 0).any()


def test_host_padding_and_schedule_mapping(tmp_path, mtx_dir):
    """chunkSize pads the height by a FULL extra chunk even when it divides (src/sparse_matrix.cpp:147-150, Q7);
    the Lift schedule name/properties select the CUDA schedule (host/inc/kernel_config.h)."""
    mtx = mtx_dir["Trefethen_20"]  # 20 x 20
    cases = {  # kernel -> (padded rows, variant, lanes, flags)
        "glb-sdp": (20, 1, 0, 4), "glb-sdp-rsa": (20, 1, 0, 0), "swrg-slcl-pmdp": (20, 2, 32, 0),
        "swrg-slcl-sdp-chunk-128": (128, 1, 0, 4), "awrg-alcl-alcl-edp-split-8": (20, 3, 0, 0),
        "awrg-alcl-fdp-chunk-rsa-8": (24, 0, 0, 0)}
    for k, (rows, variant, lanes, flags) in cases.items():
        d = _parse(tmp_path, mtx, os.path.join(KERNELS, "spmv", k + ".json"))
        i = d["info"]
        assert (i["rows"], i["variant"], i["lanes"], i["flags"]) == (rows, variant, lanes, flags), (k, i)
        assert len(d["rp"]) == rows + 1 and len(d["x"]) == rows and len(d["y"]) == rows
        assert i["output_bytes"] == 4 * rows and d["rp"][-1] == d["rp"][20] == 158
    d = _parse(tmp_path, mtx_dir["1138_bus"], os.path.join(KERNELS, "spmv", "awrg-alcl-fdp-chunk-rsa-8.json"))
    assert d["info"]["rows"] == 1144  # C1 height padded to 1144
    for alg, sr in (("spmv", 0), ("sssp", 1), ("bfs", 2), ("scc", 3)):
        elem = "float" if sr < 2 else "int"
        assert _parse(tmp_path, mtx, os.path.join(KERNELS, alg, "glb-sdp.json"), elem=elem)["info"]["semiring"] == sr


def test_host_reads_the_reference_kernel_jsons_when_present(tmp_path, mtx_dir):
    ref = "/root/reference/example"
    if not os.path.isdir(ref):
        pytest.skip("reference not present")
    want = {"spmv": 0, "pr": 0, "sssp": 1, "bfs": 2, "scc": 3}
    for alg, sr in want.items():
        for k in ("kernel", "kernel2", "kernel3", "kernel4", "kernel5", "kernel6", "kernel7"):
            d = _parse(tmp_path, mtx_dir["Trefethen_20"], f"{ref}/{alg}/{k}.json", elem="float" if sr < 2 else "int")
            assert d["info"]["semiring"] == sr, (alg, k)
            assert "Kernel: " in d["stdout"] and "param var: MHeight" in d["stdout"]


def test_host_cli_contract(tmp_path, mtx_dir):
    """Required flags exit(-1) with the reference's message; unknown flags print usage and exit(0); non-square
    input exits 2 (inc/options.h:97-105,210-248, inc/common.h:49-52)."""
    exe = os.path.join(BIN, "spmv_harness")
    r = subprocess.run([exe, "-m", mtx_dir["Trefethen_20"]], capture_output=True, text=True)
    assert r.returncode == 255 and 'Required argument "matrix_name" not set.' in r.stdout
    r = subprocess.run([exe, "--bogus"], capture_output=True, text=True)
    assert r.returncode == 0 and "Error: Invalid argument '--bogus'." in r.stdout and "Usage:" in r.stdout
    r = subprocess.run([exe, "-h"], capture_output=True, text=True)
    assert r.returncode == 0 and "--matrix_name" in r.stdout and "--timeout" in r.stdout
    rect = tmp_path / "rect.mtx"
    rect.write_text("%%MatrixMarket matrix coordinate real general\n% c\n2 3 1\n1 1 1.0\n")
    r = subprocess.run([exe, "-e", "e", "-n", "h", "-m", str(rect), "-f", "r", "-k",
                        os.path.join(KERNELS, "spmv", "glb-sdp.json"), "-r", os.path.join(KERNELS, "runfile.csv")],
                       capture_output=True, text=True, env=dict(os.environ, SHB_QUIET_LOG="1"))
    assert r.returncode == 2 and "Matrix is not square. Failing computation." in r.stdout
    cplx = tmp_path / "c.mtx"
    cplx.write_text("%%MatrixMarket matrix coordinate complex general\n1 1 1\n1 1 1.0 0.0\n")
    r = subprocess.run([os.path.join(BIN, "just_parser_harness"), "-m", str(cplx), "-k",
                        os.path.join(KERNELS, "spmv", "glb-sdp.json")], capture_output=True, text=True)
    assert r.returncode == 255 and "Cannot process this matrix type" in r.stderr


@pytest.mark.parametrize("kind", ["real general", "real symmetric", "pattern general", "integer general", "irregular"])
def test_parallel_parse_equals_sequential_parse(tmp_path, monkeypatch, kind):
    """the threaded Matrix Market entry parser (host/inc/sparse_matrix.h) gives exactly the
    sequential token parser's operator -- same tuples in the same order -- on general / symmetric / pattern /
    integer files with blank lines, CRLF and surplus entries; an irregular file (two entries on one line)
    falls back to the sequential parser."""
    rng = np.random.default_rng(11)
    n, nnz = 3000, 60000
    I = rng.integers(1, n + 1, nnz)
    J = rng.integers(1, n + 1, nnz)
    V = rng.normal(0, 40, nnz)
    field, sym = ("real", "general") if kind == "irregular" else kind.split()
    if sym == "symmetric":
        I, J = np.maximum(I, J), np.minimum(I, J)
    lines = [f"%%MatrixMarket matrix coordinate {field} {sym}", "% comment", f"{n} {n} {nnz}"]
    for k in range(nnz):
        if field == "pattern":
            e = f"{I[k]} {J[k]}"
        elif field == "integer":
            e = f"{I[k]}\t{J[k]} {int(V[k])}"
        else:
            e = f"{I[k]} {J[k]}  {V[k]:.9g}" if k % 3 else f" {I[k]} {J[k]} {V[k]:.6e} "
        lines.append(e + ("\r" if k % 50 == 0 else ""))
        if k % 1000 == 999:
            lines.append("")
    if kind == "irregular":
        lines[5] = lines[5].rstrip("\r") + " " + lines.pop(6).rstrip("\r")  # two entries on one line
    lines += ["1 1 5", "2 2 6"]  # beyond nonz: ignored by both parsers
    mtx = tmp_path / "m.mtx"
    mtx.write_text("\n".join(lines) + "\n")
    kernel = os.path.join(KERNELS, "spmv", "glb-sdp-rsa.json")
    monkeypatch.setenv("SHB_MM_THREADS", "1")
    a = _parse(tmp_path, str(mtx), kernel, extra=("--keep_values",))
    monkeypatch.setenv("SHB_MM_THREADS", "7")
    monkeypatch.setenv("SHB_MM_PARALLEL_MIN_BYTES", "1")
    b = _parse(tmp_path, str(mtx), kernel, extra=("--keep_values",))
    for k in ("rp", "col", "gold"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["val"].view(np.int32), b["val"].view(np.int32))
    assert a["rp"][-1] == (nnz if sym == "general" else nnz + int((I != J).sum()))


def test_threaded_parser_fast_tokens_match_the_c_library_on_tricky_spellings(tmp_path, monkeypatch):
    """The threaded loader's own integer / decimal token parsers (exact fast path, C library for the rest) give the
    sequential strtol / strtod tuples bit for bit: short and 15-20 digit mantissas, exponents at and past +-22,
    leading '+', '.5', '5.', '-0', hexadecimal floats, overflow to infinity, subnormal results."""
    rng = np.random.default_rng(99)
    specials = ["0", "-0", "+3", ".5", "5.", "-.25e1", "1e5", "1E-5", "1e22", "1e23", "1e-22", "1e-23", "9007199254740992",
                "9007199254740993", "12345678901234567890", "0.1234567890123456789", "123456789012345678e-10",
                "4.9e-324", "1e-320", "1e39", "3.4028235e38", "0x1.8p3", "0X10", "1e400", "-1e400", "007", "1.0000000000000002",
                "8.5e+0007", "2.5E+02", "0.000000000000000000001", "179769313486231570000", "1.7976931348623157e308"]
    vals = list(specials)
    for _ in range(30000):
        k = rng.integers(0, 6)
        m = rng.integers(0, 10 ** int(rng.integers(1, 19)))
        if k == 0:
            vals.append(str(m))
        elif k == 1:
            s = str(m)
            p = int(rng.integers(0, len(s) + 1))
            vals.append(("-" if rng.integers(2) else "") + s[:p] + "." + s[p:])
        elif k == 2:
            vals.append(f"{m}e{int(rng.integers(-30, 31))}")
        elif k == 3:
            vals.append(f"{rng.normal(0, 1e3):.17g}")
        elif k == 4:
            vals.append(f"{rng.normal(0, 1):.6e}")
        else:
            vals.append(f"{float(rng.integers(-10**6, 10**6)) / 1000:.3f}")
    n = 500
    nnz = len(vals)
    I = rng.integers(1, n + 1, nnz)
    J = rng.integers(1, n + 1, nnz)
    lines = ["%%MatrixMarket matrix coordinate real general", f"{n} {n} {nnz}"] + [f"{I[k]} {J[k]} {vals[k]}" for k in range(nnz)]
    mtx = tmp_path / "t.mtx"
    mtx.write_text("\n".join(lines) + "\n")
    kernel = os.path.join(KERNELS, "spmv", "glb-sdp-rsa.json")
    monkeypatch.setenv("SHB_MM_THREADS", "1")
    a = _parse(tmp_path, str(mtx), kernel, extra=("--keep_values",))
    monkeypatch.setenv("SHB_MM_THREADS", "5")
    monkeypatch.setenv("SHB_MM_PARALLEL_MIN_BYTES", "1")
    b = _parse(tmp_path, str(mtx), kernel, extra=("--keep_values",))
    assert a["rp"][-1] == nnz
    for k in ("rp", "col"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["val"].view(np.int32), b["val"].view(np.int32))


@pytest.mark.parametrize("name", ["1138_bus", "ca-AstroPh"])
def test_threaded_csr_encode_equals_sequential_on_padded_heights(tmp_path, mtx_dir, monkeypatch, name):
    """The CSR encoder's row-range threads (host/inc/sparse_matrix.h csr_encode) give the sequential counting sort's
    arrays on a symmetric fixture, with and without the padded rows of a chunked kernel, for thread counts that do
    and do not divide the height (and more threads than a small height leaves rows for)."""
    mtx = mtx_dir[name]
    for kernel in ("glb-sdp-rsa", "awrg-alcl-fdp-chunk-rsa-8", "swrg-slcl-sdp-chunk-128"):
        k = os.path.join(KERNELS, "spmv", kernel + ".json")
        monkeypatch.setenv("SHB_MM_THREADS", "1")
        monkeypatch.delenv("SHB_MM_PARALLEL_MIN_BYTES", raising=False)
        a = _parse(tmp_path, mtx, k)
        for threads in ("2", "7", "16"):
            monkeypatch.setenv("SHB_MM_THREADS", threads)
            monkeypatch.setenv("SHB_MM_PARALLEL_MIN_BYTES", "1")
            b = _parse(tmp_path, mtx, k)
            for key in ("rp", "col", "gold"):
                assert np.array_equal(a[key], b[key]), (kernel, threads, key)
            assert np.array_equal(a["val"].view(np.int32), b["val"].view(np.int32)), (kernel, threads)
