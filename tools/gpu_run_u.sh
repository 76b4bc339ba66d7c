#!/bin/bash
# end-of-round run: reference harness mirror (with gap filling), whole GPU suite, driver rehearsal on one GPU,
# launch list of the default bench command
mkdir -p gpurun_out
timeout 500 python tools/reference_suite.py > gpurun_out/u_reference_suite.txt 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/u_reference_suite.txt
timeout 1300 python -m pytest tests -x -q -m gpu > gpurun_out/u_full_tests.log 2>&1; echo "tests rc=$?"; tail -n 2 gpurun_out/u_full_tests.log
bash tools/gpu_run_k.sh 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/u_launches.csv \
  python bench.py --no-cpu --steps 2 --warmup 1 --points 2e8 > gpurun_out/u_launches.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/u_launches.csv
