#!/bin/bash
# The GPU suite against the bounds-checked build (csrc/Makefile `check`), preceded by a negative
# control when build/neg holds a deliberately broken build (the check must trap on it).
cd /root/repo
mkdir -p gpurun_out
if [ -f build/neg/pangeo-pyinterp_b200/csrc/libpyinterp_b200_check.so ]; then
  PYINTERP_B200_LIB=$PWD/build/neg/pangeo-pyinterp_b200/csrc/libpyinterp_b200_check.so timeout 600 \
    python -m pytest tests/test_gpu_fullsize.py -x -q -m gpu -k "odd_ny and weights" > gpurun_out/check_negative.log 2>&1
  echo "negative control (must fail): $?"
  grep -m 3 "bounds check" gpurun_out/check_negative.log
fi
PYINTERP_B200_LIB=$PWD/pangeo-pyinterp_b200/csrc/libpyinterp_b200_check.so timeout 2400 \
  python -m pytest tests -x -q -m gpu > gpurun_out/check_tests.log 2>&1
echo "checked build: $?"
tail -n 3 gpurun_out/check_tests.log
grep -c "bounds check" gpurun_out/check_tests.log
