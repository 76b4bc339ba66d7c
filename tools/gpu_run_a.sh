#!/bin/bash
# round 2, GPU call A: tests, K1 variant sweep, full bench line, launch list
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/a_smi.txt
nproc >> gpurun_out/a_smi.txt; free -g >> gpurun_out/a_smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/a_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/a_tests.log
tail -5 gpurun_out/a_tests.log
timeout 600 python tools/sweep_k1.py 1e9 > gpurun_out/a_sweep.txt 2>&1; echo "sweep rc=$?"
cat gpurun_out/a_sweep.txt
timeout 900 python bench.py > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; echo "bench rc=$?"
tail -c 6000 gpurun_out/a_bench.json; tail -5 gpurun_out/a_bench.err
