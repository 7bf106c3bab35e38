#!/bin/bash
# shb_spmv_stream_static: its test alone in a fresh process, then the whole GPU suite (the test runs last)
mkdir -p gpurun_out
echo "== alone"; timeout 400 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 300 2>&1 | tail -8
echo "== whole suite"; timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 2>&1 | tail -8
