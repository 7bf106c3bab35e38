"""-m gpu: the CUDA path against the reference's OWN device kernel text.

tests/golden/ref_*.npz hold, per fixture, what example/<alg>/kernel5.json:3 (glb-sdp, padded ELL, reversed
reduction order) and kernel6.json:3 (glb-sdp-rsa, ragged, forward order) produce when the unmodified kernel
sources run on the CPU over the reference's own cl_encode blobs with the apps' initial states
(oracle/ref_kernels.cpp, tests/golden/make_golden.py).  Here shb_spmv and shb_iterate are compared with those
arrays DIRECTLY -- no oracle restatement in between -- through the C ABI:
  * or-and, max-min, min-plus: every schedule, bit for bit (order-independent reductions)
  * plus-times: the SCALAR schedule in the kernel's own order bit for bit (forward == kernel6, reversed ==
    kernel5); every other schedule to 1e-5 relative (max norm), the north_star's float tolerance
  * the executeRun recurrence (app/sssp.cpp:112-153 and twins): same launch count, same final vector.
"""
import numpy as np
import pytest

import sparseharness_b200 as shb
from conftest import FIXTURES
from gpu_util import ALL_CFGS, DT
from test_oracle_pinned import KERNEL_ALGS, _gold, app_state, bits

pytestmark = pytest.mark.gpu
NAMES = list(FIXTURES.values())


@pytest.fixture(scope="module")
def ctx():
    c = shb.Context(0)
    yield c
    c.close()


def _same(got, want):
    """bit-identical; a NaN (the reference's pr recurrence overflows on some fixtures, TODO.md:3-9) must be a NaN
    at the same place -- its payload is hardware specific."""
    if got.dtype.kind != "f":
        return np.array_equal(got, want)
    nan = np.isnan(want)
    return np.array_equal(np.isnan(got), nan) and np.array_equal(bits(got)[~nan], bits(want)[~nan])


def _close(got, want, tol=1e-5):
    fin = np.isfinite(want)
    if not np.array_equal(np.isfinite(got), fin):
        return False
    scale = max(float(np.abs(want[fin]).max()) if fin.any() else 0.0, 1e-30)
    return float(np.abs(got[fin].astype(np.float64) - want[fin].astype(np.float64)).max()) <= tol * scale


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("alg", KERNEL_ALGS)
def test_spmv_equals_reference_kernel_text(ctx, orc, mtx_dir, name, alg):
    g = _gold(name)
    sr, elem, norm, x0, y0, alpha, beta = app_state(orc, mtx_dir[name], alg)
    h, w, rp, col, val = orc.mm_load_csr(mtx_dir[name], elem, normalise=norm)
    A = ctx.csr(h, w, rp, col, val)
    dt = DT[sr]
    xb, yb, out = ctx.array(x0.astype(dt)), ctx.array(y0.astype(dt)), ctx.empty(h * 4)
    for variant, flags in ((5, shb.F_REVERSED_ORDER), (6, 0)):
        want = g[f"k{variant}_{alg}_step1"]
        out.fill_u32(0xDEADBEEF)
        ctx.spmv(A, sr, xb, yb, alpha, beta, out, shb.launch_cfg(variant=shb.VARIANT_SCALAR, flags=flags))
        got = out.download(dt, h)
        assert _same(got, want), (name, alg, variant, "scalar")
        for label, cfg in ALL_CFGS:
            if "blocked" in label:
                continue  # covered with its own operator copies in test_gpu_spmv.py
            out.fill_u32(0xDEADBEEF)
            ctx.spmv(A, sr, xb, yb, alpha, beta, out, shb.launch_cfg(**cfg))
            got = out.download(dt, h)
            if sr == shb.PLUS_TIMES:
                assert _close(got, want), (name, alg, variant, label)
            else:
                assert _same(got, want), (name, alg, variant, label)


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("alg", ["sssp", "bfs", "pr", "scc"])
def test_iterate_equals_reference_recurrence(ctx, orc, mtx_dir, name, alg):
    g = _gold(name)
    sr, elem, norm, x0, y0, alpha, beta = app_state(orc, mtx_dir[name], alg)
    h, w, rp, col, val = orc.mm_load_csr(mtx_dir[name], elem, normalise=norm)
    A = ctx.csr(h, w, rp, col, val)
    dt = DT[sr]
    for variant, flags in ((5, shb.F_REVERSED_ORDER), (6, 0)):
        want = g[f"k{variant}_{alg}_final"]
        iters, terminated, cap = (int(v) for v in g[f"k{variant}_{alg}_iters"])
        cfgs = [shb.launch_cfg(variant=shb.VARIANT_SCALAR, flags=flags)]
        if sr != shb.PLUS_TIMES:
            cfgs += [None, shb.launch_cfg(variant=shb.VARIANT_STREAM)]
        for cfg in cfgs:
            for check_every in (1, 5):
                xb, yb, wb = ctx.array(x0.astype(dt)), ctx.array(y0.astype(dt)), ctx.empty(h * 4)
                wb.zero()
                res, n, _, _, ok = ctx.iterate(A, sr, xb, yb, alpha, beta, wb, delta=1e-4, max_iters=cap,
                                               check_every=check_every, cfg=cfg)
                got = res.download(dt, h)
                assert n == iters and (ok or iters == cap), (name, alg, variant, n, iters)
                assert _same(got, want), (name, alg, variant)
