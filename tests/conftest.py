import gzip
import os
import shutil
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
# reference example file -> SuiteSparse name (tests/golden/make_golden.py)
FIXTURES = {"matrix": "1138_bus", "matrix2": "ca-AstroPh", "matrix3": "Trefethen_20", "matrix4": "GD96_b",
            "matrix5": "arc130"}


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def mtx_dir(tmp_path_factory):
    """The golden .mtx.gz inputs, gunzipped once per session; returns {suitesparse_name: path}."""
    d = tmp_path_factory.mktemp("mtx")
    out = {}
    for name in FIXTURES.values():
        dst = os.path.join(d, name + ".mtx")
        with gzip.open(os.path.join(GOLDEN, "mtx", name + ".mtx.gz"), "rb") as fi, open(dst, "wb") as fo:
            shutil.copyfileobj(fi, fo)
        out[name] = dst
    return out


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle as o
    o.lib()
    return o
