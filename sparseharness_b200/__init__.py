"""sparseharness_b200 -- B200-native semiring CSR SpMV (the hot path of AdamBrouwersHarries/sparseharness).

The product is the C-ABI CUDA library lib/libshb.so (include/shb.h) plus the C++ harness host under
host/.  This Python package is only a thin ctypes view of that ABI for tests and bench.py.
"""
from .api import (  # noqa: F401
    MAX_MIN, MIN_PLUS, OR_AND, PLUS_TIMES, TERM_ABSDIFF, TERM_EQUAL, VARIANT_AUTO, VARIANT_BLOCKED, VARIANT_SCALAR, VARIANT_STREAM,
    VARIANT_VECTOR, F_NO_CACHE_HINTS, F_NO_PANELS, F_STREAM_MERGE, F_STREAM_NO_MERGE, F_STREAM_SCAN, F_STREAM_TMA, F_NO_Y_SKIP, F_REVERSED_ORDER, F_Y_IS_ONE, Buffer, Context, Csr, Group, GroupPartition, ShbError,
    device_count, launch_cfg)
