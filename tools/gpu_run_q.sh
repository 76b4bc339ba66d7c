#!/bin/bash
# K4s on float32 grids: tests, then timings
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_parity.py tests/test_gpu_golden.py -x -q -m gpu -k "bicubic or windowed or golden" > gpurun_out/q_tests.log 2>&1
echo "tests: $?"; tail -n 3 gpurun_out/q_tests.log
timeout 300 python tools/bicubic_dtype_bench.py > gpurun_out/q_bicubic.txt 2>&1
echo "bench: $?"; cat gpurun_out/q_bicubic.txt
