#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/sweep_k1.py 1e9 > gpurun_out/d_sweep.txt 2>&1; echo "sweep rc=$?"; cat gpurun_out/d_sweep.txt
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -k "bilinear or temporal or offset or seam or tolerance" > gpurun_out/d_tests.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/d_tests.log
