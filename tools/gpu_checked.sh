#!/bin/bash
# The GPU test suite against the CHECKED build of the library (device-side index checks, shb_internal.cuh): stands in
# for the compute-sanitizer sweep this pool does not allow.  A failed check prints its condition and traps the launch,
# which fails the test.  Build first: make -C sparseharness_b200/csrc CHECKED=1
mkdir -p gpurun_out
test -f sparseharness_b200/lib/libshb_checked.so || { echo "libshb_checked.so missing"; exit 1; }
SHB_LIB=checked timeout 1500 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/checked_pytest.txt 2>&1
echo "rc=$?"; tail -6 gpurun_out/checked_pytest.txt; grep -c "SHB_DEV_CHECK failed" gpurun_out/checked_pytest.txt
SHB_LIB=checked timeout 200 python tools/sanitize_driver.py > gpurun_out/checked_driver.txt 2>&1; echo "driver rc=$?"; tail -2 gpurun_out/checked_driver.txt
