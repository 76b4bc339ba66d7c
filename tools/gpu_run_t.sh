#!/bin/bash
# two GPUs: the multi-GPU tests, then the driver's round-end commands (tools/gpu_run_k.sh) at N = 2
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/t_multi_tests.log 2>&1; echo "multi tests rc=$?"; tail -3 gpurun_out/t_multi_tests.log
bash tools/gpu_run_k.sh 2
