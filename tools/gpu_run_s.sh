#!/bin/bash
# ncu after the evaluator rewrite: idw kernel (--set full), and the launch list of the default bench command
mkdir -p gpurun_out /tmp/ncu
NCU="ncu --set full --clock-control none --import-source on"
B="python bench.py --no-cpu --no-secondary --steps 2 --warmup 1"
timeout 600 $B --workload idw --points 5e7 > gpurun_out/s_idw.json 2> gpurun_out/s_idw.err && \
timeout 900 $NCU -k regex:bilinear2d_fast -s 3 -c 1 -o /tmp/ncu/s_idw -f $B --workload idw --points 5e7 > gpurun_out/s_idw_ncu.log 2>&1
echo "idw rc=$?"
ncu -i /tmp/ncu/s_idw.ncu-rep --page raw --csv > gpurun_out/s_idw_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/s_idw_raw.csv 5e7 > gpurun_out/s_idw_ncu_summary.txt 2>&1
timeout 900 python bench.py --no-cpu --steps 2 --warmup 1 --points 2e8 > gpurun_out/s_all.json 2> gpurun_out/s_all.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/s_launches.csv \
  python bench.py --no-cpu --steps 2 --warmup 1 --points 2e8 > gpurun_out/s_launches.log 2>&1
echo "launch list rc=$?"; du -sh gpurun_out
