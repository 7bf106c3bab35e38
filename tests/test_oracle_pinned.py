"""Pins the CPU oracle (oracle/shb_oracle.c) to the reference's own outputs.

tests/golden/ref_*.npz were produced by the UNMODIFIED reference code (oracle/_ref/ref_driver, see
tests/golden/make_golden.py): loader rows (src/sparse_matrix.cpp:11-119), normalisations (:409-456)
and Gold<float>::spmv (inc/spmv_gold.h:9-28).  Everything here is bit-exact.
"""
import os

import numpy as np
import pytest

from conftest import FIXTURES, GOLDEN

NAMES = list(FIXTURES.values())


def _gold(name):
    return np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))


@pytest.mark.parametrize("name", NAMES)
def test_loader_matches_reference_rows(orc, mtx_dir, name):
    g = _gold(name)
    h, w, rp, col, val = orc.mm_load_csr(mtx_dir[name], "float")
    assert (h, w) == tuple(g["shape"])
    assert np.array_equal(rp, g["f_row_ptr"]) and np.array_equal(col, g["f_col"])
    assert np.array_equal(val.view(np.int32), g["f_val"].view(np.int32))  # Q1-Q4 incl. int truncation
    _, _, rp, col, val = orc.mm_load_csr(mtx_dir[name], "int")
    assert np.array_equal(rp, g["i_row_ptr"]) and np.array_equal(col, g["i_col"]) and np.array_equal(val, g["i_val"])


@pytest.mark.parametrize("name", NAMES)
def test_normalisations_match_reference(orc, mtx_dir, name):
    g = _gold(name)
    _, _, _, _, val = orc.mm_load_csr(mtx_dir[name], "int", normalise="scc")
    assert np.array_equal(val, g["scc_val"])
    _, _, _, _, val = orc.mm_load_csr(mtx_dir[name], "float", normalise="pr")
    assert np.array_equal(val.view(np.int32), g["pr_val"].view(np.int32))


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("xmode", [0, 1])
def test_gold_and_kernel_restatement_match_reference_gold(orc, mtx_dir, name, xmode):
    g = _gold(name)
    h, w, rp, col, val = orc.mm_load_csr(mtx_dir[name], "float")
    x = orc.xmode_vector(h, xmode)
    y = np.zeros(h, np.float32)
    want = g[f"gold_x{xmode}"]
    got = orc.gold_spmv(rp, col, val, x, y)
    assert np.array_equal(got.view(np.int32), want.view(np.int32))
    # the kernel restatement (device formula), both accumulation orders, 1 and N threads
    for order in (0, 1):
        for threads in (1, 4):
            out = orc.spmv(orc.PLUS_TIMES, rp, col, val, x, y, 1.0, 0.0, order=order, threads=threads)
            assert np.array_equal(out, want), (order, threads)


def test_known_answers_from_survey(orc):
    # values obtained by running the reference; regenerated here from the npz
    g = _gold("1138_bus")["gold_x0"]
    assert list(g[:8]) == [1460, 1, 2, 2, 0, 2, 3, 1] and g.astype(np.float64).sum() == 2270
    assert _gold("ca-AstroPh")["gold_x0"][0] == 43 and _gold("ca-AstroPh")["gold_x0"].astype(np.float64).sum() == 396160
    t = _gold("Trefethen_20")["gold_x0"]
    assert t[0] == 7 and t[1] == 9 and t.sum() == 777
    assert _gold("GD96_b")["gold_x0"][0] == 47 and _gold("GD96_b")["gold_x0"].sum() == 193
    assert _gold("arc130")["gold_x0"][0] == 1 and _gold("arc130")["gold_x0"].astype(np.float64).sum() == -4717836


def test_live_reference_when_present(orc, mtx_dir):
    """In the build container the reference binary itself is compared (skipped on the GPU box)."""
    if not (orc.ref_available() and os.path.isdir("/root/reference")):
        pytest.skip("oracle/_ref/ref_driver or /root/reference not present")
    for name in ("Trefethen_20", "GD96_b"):
        h, w, rp, col, val = orc.ref_rows(mtx_dir[name], "float")
        _, _, rp2, col2, val2 = orc.mm_load_csr(mtx_dir[name], "float")
        assert np.array_equal(rp, rp2) and np.array_equal(col, col2) and np.array_equal(val, val2)
        assert np.array_equal(orc.ref_gold(mtx_dir[name], 1),
                              orc.gold_spmv(rp2, col2, val2, orc.xmode_vector(h, 1), np.zeros(h, np.float32)))


def test_sample_file_for_the_reference_arm_round_trips(orc, tmp_path):
    """bench.py --impl reference hands a CSR sample to the reference as Matrix Market text: the file written with
    second_index_is_row makes the reference's loader (which takes the second index as the row, Q1) rebuild exactly
    those rows, and its Gold agrees with the oracle on them (values int-truncated by the loader, Q2)."""
    rp, col, val = orc.gen_uniform(4096, 4096, 16, val_mode=1, row_begin=0, row_count=300)
    path = str(tmp_path / "sample.mtx")
    orc.write_mtx(path, 300, 4096, rp, col, val, second_index_is_row=True)
    head = open(path).read().splitlines()[:3]
    assert head[0].startswith("%%MatrixMarket matrix coordinate real general") and head[1] == f"300 4096 {len(col)}"
    assert head[2].split()[:2] == [str(col[0] + 1), "1"]
    if not orc.ref_available():
        pytest.skip("oracle/_ref/ref_driver not present")
    h, w, rrp, rcol, rval = orc.ref_rows(path, "float")
    assert (h, w) == (300, 4096) and np.array_equal(rrp, rp)
    for r in range(300):  # the reference keeps file order inside a row (its per-row sort works on a copy)
        a, b = rp[r], rp[r + 1]
        assert sorted(zip(rcol[a:b], rval[a:b])) == sorted(zip(col[a:b], np.trunc(val[a:b])))
    t = orc.reference_gold_time(path, 1, 2)
    assert t is not None and t[0] > 0


# ---- the reference's OWN device kernel text, executed on the CPU (oracle/ref_kernels.cpp) -------------------
# k5 = example/<alg>/kernel5.json:3 (glb-sdp: padded ELL, products parked in the temp buffer at W-1-i and reduced
# ascending => REVERSED order), k6 = kernel6.json:3 (glb-sdp-rsa: ragged, forward order); blobs from the
# reference's own cl_encode; initial state, recurrence and termination test of the apps.  Bit-exact, floats too.
def app_state(orc, mtx, alg, n=None):
    """(semiring, elem, normalise, x0, y0, alpha, beta) of app/<alg>.cpp main(); cites in oracle/ref_kernels.cpp."""
    if alg in ("spmv", "spmvx1"):
        h = orc.mm_load_csr(mtx, "float")[0]
        return orc.PLUS_TIMES, "float", "none", orc.xmode_vector(h, 0 if alg == "spmv" else 1), np.zeros(h, np.float32), 1.0, 0.0
    if alg == "sssp":
        h = orc.mm_load_csr(mtx, "float")[0]
        x0 = np.full(h, orc.FLT_MAX, np.float32); x0[0] = 0
        return orc.MIN_PLUS, "float", "none", x0, x0.copy(), 0.0, 0.0
    if alg == "bfs":
        h = orc.mm_load_csr(mtx, "int")[0]
        x0 = np.zeros(h, np.int32); x0[0] = 1
        return orc.OR_AND, "int", "none", x0, x0.copy(), 1, 0
    if alg == "pr":
        h = orc.mm_load_csr(mtx, "float")[0]
        x0 = np.full(h, np.float32(1.0) / np.float32(h), np.float32)
        beta = (np.float32(1.0) - np.float32(0.85)) / np.float32(h)
        return orc.PLUS_TIMES, "float", "pr", x0, np.ones(h, np.float32), 1.0, float(beta)
    if alg == "scc":
        h = orc.mm_load_csr(mtx, "int")[0]
        return (orc.MAX_MIN, "int", "scc", np.arange(h, dtype=np.int32), np.full(h, orc.INT_MIN, np.int32),
                int(orc.INT_MAX), int(orc.INT_MIN))
    raise ValueError(alg)


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


KERNEL_ALGS = ["spmv", "spmvx1", "sssp", "bfs", "pr", "scc"]


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("alg", KERNEL_ALGS)
@pytest.mark.parametrize("variant", [5, 6])
def test_restatement_matches_reference_kernel_text_one_launch(orc, mtx_dir, name, alg, variant):
    g = _gold(name)
    sr, elem, norm, x0, y0, alpha, beta = app_state(orc, mtx_dir[name], alg)
    h, w, rp, col, val = orc.mm_load_csr(mtx_dir[name], elem, normalise=norm)
    want = g[f"k{variant}_{alg}_step1"]
    assert len(want) == h
    for threads in (1, 3):
        got = orc.spmv(sr, rp, col, val, x0, y0, alpha, beta, order=1 if variant == 5 else 0, threads=threads)
        assert np.array_equal(bits(got), bits(want)), (name, alg, variant)


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("alg", ["sssp", "bfs", "pr", "scc"])
@pytest.mark.parametrize("variant", [5, 6])
def test_restatement_matches_reference_recurrence(orc, mtx_dir, name, alg, variant):
    g = _gold(name)
    sr, elem, norm, x0, y0, alpha, beta = app_state(orc, mtx_dir[name], alg)
    h, w, rp, col, val = orc.mm_load_csr(mtx_dir[name], elem, normalise=norm)
    want = g[f"k{variant}_{alg}_final"]
    iters, terminated, cap = (int(v) for v in g[f"k{variant}_{alg}_iters"])
    got, n = orc.iterate(sr, rp, col, val, x0, y0, alpha, beta, delta=1e-4, max_iters=cap,
                         order=1 if variant == 5 else 0)
    assert n == iters and (terminated == 1 or iters == cap)
    assert np.array_equal(bits(got), bits(want)), (name, alg, variant)
