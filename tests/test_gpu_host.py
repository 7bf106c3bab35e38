"""GPU tests of the C++ harness executables (host/bin/*_harness): the reference's CLI, Matrix Market input,
kernel-JSON and runfile in, the reference's result rows out, computed by libshb on the device and compared
with the oracle (faithful recurrences incl. padded rows; canonical BFS / SSSP / PageRank answers)."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import FIXTURES, GOLDEN, ROOT

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "host", "bin")
KERNELS = os.path.join(GOLDEN, "kernels")
ROW = re.compile(r'\(([-0-9.e+]+), "(\w+)", "([^"]*)", (\d+), (\d+), "([^"]*)", "([^"]*)", "([^"]*)",(\d+),(\d+),"(\w+)", "([^"]*)"\)')


def run_app(app, mtx, kernel, tmp_path, *extra, runfile="runfile.csv", trials=3, name="m"):
    dump = str(tmp_path / f"{app}.bin")
    cmd = [os.path.join(BIN, f"{app}_harness"), "-e", "exp1", "-n", "testhost", "-m", mtx, "-f", name, "-k", kernel, "-d",
           "0", "-r", os.path.join(KERNELS, runfile), "-i", str(trials), "-t", "2000", "--dump", dump, *extra]
    r = subprocess.run(cmd, capture_output=True, text=True, env=dict(os.environ, SHB_QUIET_LOG="1"), timeout=600)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    inserts = [l for l in r.stdout.splitlines() if l.startswith("INSERT INTO table_name")]
    return r, inserts, dump


def pad(rp, n_pad):
    return np.concatenate([rp, np.full(n_pad - (len(rp) - 1), rp[-1], rp.dtype)])


@pytest.mark.parametrize("kernel", ["awrg-alcl-fdp-chunk-rsa-8", "glb-sdp", "glb-sdp-rsa", "swrg-slcl-pmdp",
                                    "swrg-slcl-sdp-chunk-128", "awrg-alcl-alcl-edp-split-8"])
def test_spmv_harness_config1_reports_correct(tmp_path, mtx_dir, kernel):
    """BASELINE config 1: example/matrix.mtx (HB/1138_bus), x = 1, validated against the in-run gold."""
    r, inserts, dump = run_app("spmv", mtx_dir["1138_bus"], os.path.join(KERNELS, "spmv", kernel + ".json"), tmp_path,
                               trials=10, name="matrix3")
    assert len(inserts) == 1
    assert inserts[0].startswith("INSERT INTO table_name (time, correct, kernel, global, local, host, device, matrix, "
                                 "iteration, trial,statistic, experiment_id) VALUES (") and inserts[0].endswith(");")
    rows = ROW.findall(inserts[0])
    raw = [x for x in rows if x[10] == "RAW_RESULT"]
    assert 1 <= len(raw) <= 10 and all(x[1] == "correct" for x in raw), rows
    assert rows[-1][10] == "MEDIAN_RESULT" and rows[-1][1] == "statisticvalue"
    assert all(x[2] == kernel and x[3] == "1280" and x[4] == "1" and x[5] == "testhost" and x[7] == "matrix3" and
               x[11] == "exp1" and "B200" in x[6] for x in rows)
    assert "Benchmarking run: {1280 1 1 / 1 1 1}" in r.stdout and "runtimes: [" in r.stdout
    assert 'PROFILING_DATUM("clEnqueueNDRangeKernel", "harness", ' in r.stdout
    assert "kernel has probably failed" not in r.stderr
    g = np.load(os.path.join(GOLDEN, "ref_1138_bus.npz"))["gold_x0"]
    assert np.array_equal(np.fromfile(dump, np.float32), g)


@pytest.mark.parametrize("name", list(FIXTURES.values()))
def test_spmv_harness_all_fixtures_two_runs(tmp_path, mtx_dir, name):
    r, inserts, dump = run_app("spmv", mtx_dir[name], os.path.join(KERNELS, "spmv", "awrg-alcl-fdp-chunk-rsa-8.json"),
                               tmp_path, runfile="runfile2.csv")
    assert len(inserts) == 2
    for ins in inserts:
        assert all(x[1] in ("correct", "statisticvalue") for x in ROW.findall(ins))
    g = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))["gold_x0"]
    assert np.array_equal(np.fromfile(dump, np.float32).view(np.int32), g.view(np.int32))


def faithful_case(orc, app, mtx, tmp_path, kernel, cap=100000, host_loop=False):
    elem = "float" if app in ("sssp", "pr") else "int"
    norm = {"pr": "pr", "scc": "scc"}.get(app, "none")
    h, w, rp, col, val = orc.mm_load_csr(mtx, elem, normalise=norm)
    chunk = 8 if "chunk-rsa-8" in kernel else 128 if "chunk-128" in kernel else 0
    hp = h + (chunk - h % chunk) if chunk else h
    rp = pad(rp, hp)
    i = np.arange(hp)
    if app == "sssp":
        sr, x0, alpha, beta = orc.MIN_PLUS, np.where(i == 0, 0, orc.FLT_MAX).astype(np.float32), 0.0, 0.0
        y0 = x0
    elif app == "bfs":
        sr, x0, alpha, beta = orc.OR_AND, (i == 0).astype(np.int32), 1, 0
        y0 = x0
    elif app == "pr":
        sr, x0 = orc.PLUS_TIMES, np.full(hp, np.float32(1.0) / np.float32(h), np.float32)
        y0, alpha, beta = np.ones(hp, np.float32), 1.0, (np.float32(1.0) - np.float32(0.85)) / np.float32(h)
    else:
        sr, x0, y0, alpha, beta = orc.MAX_MIN, i.astype(np.int32), np.full(hp, orc.INT_MIN, np.int32), orc.INT_MAX, orc.INT_MIN
    want, launches = orc.iterate(sr, rp, col, val, x0, y0, alpha, beta, delta=1e-4, y_rebind=True, max_iters=cap)
    alg_dir = "spmv" if app == "pr" else app
    extra = ["--max_iterations", str(cap)] + (["--host_loop"] if host_loop else [])
    r, inserts, dump = run_app(app, mtx, os.path.join(KERNELS, alg_dir, kernel + ".json"), tmp_path, *extra, trials=2)
    got = np.fromfile(dump, want.dtype)
    assert len(inserts) == 2  # one INSERT per trial
    for t, ins in enumerate(inserts):
        rows = ROW.findall(ins)
        raw = [x for x in rows if x[10] == "RAW_RESULT"]
        assert len(raw) == launches, (app, kernel, len(raw), launches)
        assert sorted(int(x[8]) for x in raw) == list(range(launches)) and all(int(x[9]) == t for x in raw)
        assert all(x[1] == "notchecked" for x in rows)
        assert [x[10] for x in rows[-2:]] == ["MEDIAN_RESULT", "MULTI_ITERATION_SUM"]
        times = [float(x[0]) for x in rows]
        assert abs(times[-1] - sum(times[:-1])) <= 1e-3 * max(times[-1], 1e-9) + 1e-5  # sum includes the median row
    want = want[:h]
    if want.dtype == np.float32:
        # NaNs (1138_bus has zero column sums, so its faithful PageRank operator holds NaN / garbage like the
        # reference's) compare as NaNs: CPU and GPU produce different NaN payloads
        nan = np.isnan(want)
        assert np.array_equal(np.isnan(got), nan), (app, kernel)
        got, want = np.where(nan, np.float32(0), got), np.where(nan, np.float32(0), want)
    assert np.array_equal(got.view(np.int32), want.view(np.int32)), (app, kernel)
    return launches


@pytest.mark.parametrize("host_loop", [False, True])
def test_iterative_harnesses_faithful_mode_match_oracle(tmp_path, mtx_dir, orc, host_loop):
    t20, bus, gd = mtx_dir["Trefethen_20"], mtx_dir["1138_bus"], mtx_dir["GD96_b"]
    assert faithful_case(orc, "sssp", t20, tmp_path, "glb-sdp-rsa", host_loop=host_loop) >= 2
    assert faithful_case(orc, "sssp", bus, tmp_path, "awrg-alcl-fdp-chunk-rsa-8", host_loop=host_loop) >= 2
    faithful_case(orc, "scc", t20, tmp_path, "swrg-slcl-sdp-chunk-128", host_loop=host_loop)
    faithful_case(orc, "scc", gd, tmp_path, "glb-sdp", host_loop=host_loop)
    faithful_case(orc, "pr", bus, tmp_path, "awrg-alcl-fdp-chunk-rsa-8", host_loop=host_loop)
    faithful_case(orc, "pr", t20, tmp_path, "swrg-slcl-pmdp", host_loop=host_loop)
    # the reference's bfs recurrence may never settle (Q8): the cap ends it, result = the oracle's at the cap
    faithful_case(orc, "bfs", bus, tmp_path, "awrg-alcl-alcl-edp-split-8", cap=40, host_loop=host_loop)
    faithful_case(orc, "bfs", gd, tmp_path, "glb-sdp-rsa", cap=40, host_loop=host_loop)


def test_canonical_modes_match_oracle(tmp_path, mtx_dir, orc):
    bus = mtx_dir["1138_bus"]
    # BFS levels (edges whose value truncates to 0 are no edges, as in bool_and)
    h, w, rp, col, val = orc.mm_load_csr(bus, "int")
    want, nlev, _ = orc.bfs_levels(rp, col, val, 0)
    r, ins, dump = run_app("bfs", bus, os.path.join(KERNELS, "bfs", "glb-sdp-rsa.json"), tmp_path, "--mode", "canonical")
    assert np.array_equal(np.fromfile(dump, np.int32), want) and "BFS: iterations" in r.stdout
    rows = ROW.findall(ins[0])
    assert [x[10] for x in rows] == ["RAW_RESULT"] * 3 + ["MEDIAN_RESULT"]
    # SSSP with the file's float weights kept
    h, w, rp, col, val = orc.mm_load_csr(bus, "float", truncate=False)
    want = orc.sssp(rp, col, val, 0)
    r, ins, dump = run_app("sssp", bus, os.path.join(KERNELS, "sssp", "glb-sdp-rsa.json"), tmp_path, "--mode", "canonical")
    assert np.array_equal(np.fromfile(dump, np.float32).view(np.int32), want.view(np.int32))
    # PageRank: column-stochastic float operator, y == 1
    h, w, rp, col, val = orc.mm_load_csr(mtx_dir["GD96_b"], "float", normalise="pr", truncate=False)
    want, iters = orc.pagerank(rp, col, val, delta=1e-6, max_iters=500)
    r, ins, dump = run_app("pr", mtx_dir["GD96_b"], os.path.join(KERNELS, "spmv", "glb-sdp-rsa.json"), tmp_path, "--mode",
                           "canonical", "-c", "1e-6", "--max_iterations", "500")
    got = np.fromfile(dump, np.float32)
    assert np.abs(got - want).max() <= 1e-5 * np.abs(want).max()  # tolerance stated by BASELINE.json north_star
    assert abs(int(ROW.findall(ins[0])[0][8]) - iters) <= 1


def test_harness_runs_synthetic_configs(tmp_path, mtx_dir):
    """--synthetic builds BASELINE's synthetic operators on the device behind the same CLI (small sizes here)."""
    k = os.path.join(KERNELS, "spmv", "awrg-alcl-fdp-chunk-rsa-8.json")
    r, ins, dump = run_app("spmv", "none", k, tmp_path, "--synthetic", "uniform-int:65536:16")
    out = np.fromfile(dump, np.float32)
    assert len(out) == 65536 and all(x[1] in ("notchecked", "statisticvalue") for x in ROW.findall(ins[0]))
    from oracle import oracle as orc
    rp, col, val = orc.gen_uniform(65536, 65536, 16, val_mode=1)
    want = orc.spmv(orc.PLUS_TIMES, rp, col, val, np.ones(65536, np.float32), np.zeros(65536, np.float32), 1.0, 0.0)
    assert np.array_equal(out, want)
    r, ins, dump = run_app("sssp", "none", os.path.join(KERNELS, "sssp", "glb-sdp-rsa.json"), tmp_path, "--synthetic",
                           "grid:96:64", "--mode", "canonical")
    grp, gcol, gval = orc.gen_grid2d(96, 64)
    assert np.array_equal(np.fromfile(dump, np.float32), orc.sssp(grp, gcol, gval, 0))
    r, ins, dump = run_app("bfs", "none", os.path.join(KERNELS, "bfs", "glb-sdp-rsa.json"), tmp_path, "--synthetic",
                           "rmat:12:8", "--mode", "canonical")
    rp, col = orc.gen_rmat(12, 8)
    assert np.array_equal(np.fromfile(dump, np.int32), orc.bfs_levels(rp, col, None, 0)[0])


def test_sweep_scripts_emit_the_reference_result_files(tmp_path, mtx_dir):
    """host/scripts/experiments/run_all.sh sweeps (matrix x kernel JSON) with spmv_harness and
    leaves results/results-<id>/<matrix>/result_<kernel>.txt holding the reference's INSERT INTO / PROFILING_DATUM
    lines; gather_all.sh aggregates them.  Every (matrix, kernel) pair must come back "correct"."""
    import shutil
    data = tmp_path / "data"
    names = ["Trefethen_20", "GD96_b"]
    for n in names:
        (data / n).mkdir(parents=True)
        shutil.copy(mtx_dir[n], data / n / f"{n}.mtx")
    (data / "datasets.txt").write_text("\n".join(names) + "\n")
    scripts = os.path.join(ROOT, "host", "scripts", "experiments")
    kernels = os.path.join(KERNELS, "spmv")
    env = dict(os.environ, TRIALS="2", TIMEOUT_MS="2000", SHB_QUIET_LOG="1")
    r = subprocess.run(["bash", os.path.join(scripts, "run_all.sh"), str(data), os.path.join(BIN, "spmv_harness"), kernels,
                        os.path.join(KERNELS, "runfile.csv"), "0", "0", str(tmp_path / "scratch")], cwd=tmp_path,
                       capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    res = (tmp_path / ".last_results").read_text().strip()
    n_kernels = len([k for k in os.listdir(kernels) if k.endswith(".json")])
    files = sorted((tmp_path / res).glob("*/result_*.txt"))
    assert len(files) == len(names) * n_kernels
    assert all("INSERT INTO" in f.read_text() for f in files)
    g = subprocess.run(["sh", os.path.join(scripts, "gather_all.sh"), res], cwd=tmp_path, capture_output=True, text=True)
    assert g.returncode == 0, g.stderr
    rows = (tmp_path / "aggregate_summary.csv").read_text().splitlines()[1:]
    raw = [x.split(",") for x in rows if ",RAW_RESULT," in x]
    assert len(raw) == len(names) * n_kernels * 2 and all(x[3] == "correct" for x in raw), raw[:3]  # 2 trials, 1 run
    assert {x[0] for x in raw} == set(names)
    assert sum(",MEDIAN_RESULT," in x for x in rows) == len(names) * n_kernels
    assert "PROFILING_DATUM" in (tmp_path / "aggregate_profiling_data.txt").read_text()


@pytest.mark.parametrize("name", ["1138_bus", "GD96_b", "arc130", "ca-AstroPh", "Trefethen_20"])
def test_device_encode_equals_host_encode_and_reference_rows(tmp_path, mtx_dir, name):
    """COO -> CSR on the device (shb_csr_from_coo, --gpu_encode) builds exactly the rows the
    reference's loader builds (tests/golden/ref_*.npz, Q1-Q4: row = second index, file order inside a row, values
    through int) -- float and int operators, chunked kernels' padded heights and both normalisations included."""
    from test_host_cpu import KERNELS, _parse
    g = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))
    d = _parse(tmp_path, mtx_dir[name], os.path.join(KERNELS, "spmv", "glb-sdp-rsa.json"), extra=["--gpu_encode"])
    assert np.array_equal(d["rp"], g["f_row_ptr"]) and np.array_equal(d["col"], g["f_col"])
    assert np.array_equal(d["val"].view(np.int32), g["f_val"].view(np.int32))
    d = _parse(tmp_path, mtx_dir[name], os.path.join(KERNELS, "bfs", "glb-sdp.json"), elem="int", extra=["--gpu_encode"])
    assert np.array_equal(d["rp"], g["i_row_ptr"]) and np.array_equal(d["col"], g["i_col"]) and np.array_equal(d["val"], g["i_val"])
    for elem, norm, kern in (("int", "scc", "scc/swrg-slcl-sdp-chunk-128.json"), ("float", "pr", "pr/glb-sdp.json"),
                             ("float", "none", "spmv/awrg-alcl-fdp-chunk-rsa-8.json")):
        path = os.path.join(KERNELS, kern)
        if not os.path.exists(path):
            path = os.path.join(KERNELS, "spmv", os.path.basename(kern))
        host = _parse(tmp_path, mtx_dir[name], path, elem=elem, normalise=norm)
        dev = _parse(tmp_path, mtx_dir[name], path, elem=elem, normalise=norm, extra=["--gpu_encode"])
        assert np.array_equal(host["rp"], dev["rp"]) and np.array_equal(host["col"], dev["col"]), (elem, norm)
        assert np.array_equal(host["val"].view(np.int32), dev["val"].view(np.int32)), (elem, norm)
    # keep_values (canonical operators): no truncation on either side
    host = _parse(tmp_path, mtx_dir[name], os.path.join(KERNELS, "spmv", "glb-sdp.json"), extra=["--keep_values"])
    dev = _parse(tmp_path, mtx_dir[name], os.path.join(KERNELS, "spmv", "glb-sdp.json"), extra=["--keep_values", "--gpu_encode"])
    assert np.array_equal(host["val"].view(np.int32), dev["val"].view(np.int32)) and np.array_equal(host["col"], dev["col"])


def test_spmv_harness_with_device_encode_reports_correct(tmp_path, mtx_dir):
    """config 1 with --gpu_encode: the rows the device builds from the tuples give the gold's values ("correct")."""
    r, inserts, dump = run_app("spmv", mtx_dir["1138_bus"], os.path.join(KERNELS, "spmv", "awrg-alcl-fdp-chunk-rsa-8.json"), tmp_path,
                               "--gpu_encode", trials=2)
    rows = ROW.findall("\n".join(inserts))
    assert rows and all(row[1] in ("correct", "statisticvalue") for row in rows) and any(row[1] == "correct" for row in rows)


def test_eigenvector_harness_faithful_and_canonical(tmp_path, mtx_dir, orc):
    """eigenvector_harness (dead code in the reference, app/eigenvector.cpp:42-137): faithful = x <- A x capped at 10
    launches; canonical = normalised power iteration.  Both against the oracle."""
    kern = os.path.join(KERNELS, "spmv", "glb-sdp-rsa.json")
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir["Trefethen_20"], "float")
    x0 = np.ones(n, np.float32)
    r, inserts, dump = run_app("eigenvector", mtx_dir["Trefethen_20"], kern, tmp_path, trials=1)
    want, wit = orc.iterate(orc.PLUS_TIMES, rp, col, val, x0, np.zeros(n, np.float32), 1.0, 0.0, delta=1e-4, max_iters=10)
    got = np.fromfile(dump, np.float32)
    assert wit == 10 and "Iterations: 10" in r.stdout and np.array_equal(got[:n].view(np.uint32), want.view(np.uint32))
    n, _, rp, col, val = orc.mm_load_csr(mtx_dir["Trefethen_20"], "float", truncate=False)
    r, inserts, dump = run_app("eigenvector", mtx_dir["Trefethen_20"], kern, tmp_path, "--mode", "canonical", "--variant", "scalar",
                               "--max_iterations", "200", "-c", "1e-6", trials=1)
    want, wev, wit = orc.power_iteration(rp, col, val, x0, delta=1e-6, max_iters=200)
    got = np.fromfile(dump, np.float32)
    assert f"iterations {wit}," in r.stdout and np.array_equal(got[:n].view(np.uint32), want.view(np.uint32))
    assert "Dominant eigenvalue estimate:" in r.stdout
