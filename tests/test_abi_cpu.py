"""CPU-side checks of the C-ABI boundary: libshb.so loads, exports every symbol include/shb.h declares,
and refuses to run without a CUDA device (no CPU fallback)."""
import ctypes
import os
import re
import subprocess

import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "shb.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(?:int|const char \*)\s*(shb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from sparseharness_b200 import _capi
    names = _declared()
    assert len(names) >= 40
    for n in names:
        assert hasattr(_capi.lib, n), f"libshb.so does not export {n}"
        assert n in _capi.SIGNATURES, f"{n} has no ctypes signature"
    assert sorted(_capi.SIGNATURES) == names  # the binding declares nothing the header does not
    assert _capi.lib.shb_abi_version() == 1


def test_only_sm100a_code_in_library():
    from sparseharness_b200 import _capi
    out = subprocess.run(["cuobjdump", "-lelf", _capi.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump not available")
    archs = set(re.findall(r"sm_\d+a?", out.stdout))
    assert archs == {"sm_100a"}, archs


def test_error_strings_and_no_device_behaviour():
    import sparseharness_b200 as shb
    from sparseharness_b200 import _capi
    assert _capi.lib.shb_status_string(0) == b"SHB_OK"
    assert _capi.lib.shb_status_string(3) == b"SHB_ERR_NO_DEVICE"
    if shb.device_count() == 0:
        with pytest.raises(shb.ShbError) as e:
            shb.Context(0)
        assert e.value.status == 3 and "No CUDA devices found" in str(e.value)
        n = ctypes.c_uint64(0)
        assert _capi.lib.shb_device_max_alloc_bytes(0, ctypes.byref(n)) == 3


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under sparseharness_b200/ or host/ may reference it."""
    bad = []
    for base in ("sparseharness_b200", "host"):
        for dp, _, fns in os.walk(os.path.join(ROOT, base)):
            if "_obj" in dp or "_build" in dp:
                continue
            for fn in fns:
                if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")) or fn == "Makefile":
                    txt = open(os.path.join(dp, fn), errors="ignore").read()
                    if re.search(r"(from|import)\s+oracle|oracle/|liboracle|orc_", txt):
                        # comments that merely NAME the oracle file for cross-reference are allowed in .cu
                        hits = [l for l in txt.splitlines() if re.search(r"(from|import)\s+oracle|liboracle|orc_[a-z]+\s*\(", l)
                                and not l.strip().startswith(("//", "*", "#"))]
                        if hits:
                            bad.append((fn, hits[:2]))
    assert not bad, bad
