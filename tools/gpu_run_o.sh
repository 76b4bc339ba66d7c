#!/bin/bash
# descending-axis fast path: full GPU suite, then the default bench line (regression check)
cd /root/repo
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/o_tests.log 2>&1
echo "tests: $?"
tail -n 5 gpurun_out/o_tests.log
timeout 900 python bench.py > gpurun_out/o_bench.json 2> gpurun_out/o_bench.err
echo "bench: $?"
timeout 300 python tools/desc_bench.py > gpurun_out/o_desc.txt 2>&1
echo "desc: $?"
cat gpurun_out/o_desc.txt
