#!/bin/bash
# the allocation-free loess entry point: alone in a fresh process, then after every other GPU test
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/n_alone.log 2>&1
echo "alone: $?" | tee -a gpurun_out/n_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/n_all.log 2>&1
echo "all: $?" | tee -a gpurun_out/n_all.log
tail -3 gpurun_out/n_alone.log gpurun_out/n_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/n_smoke.log 2>&1
echo "smoke: $?"
