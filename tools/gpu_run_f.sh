#!/bin/bash
mkdir -p gpurun_out /tmp/ncu
NCU="ncu --set full --clock-control none --import-source on"
B="python bench.py --no-cpu --no-secondary --steps 2 --warmup 1 --workload bicubic --points 5e7"
timeout 300 $B > gpurun_out/f_bicubic.json 2> gpurun_out/f_bicubic.err && \
timeout 900 $NCU -k regex:bicubic_sep -s 3 -c 1 -o /tmp/ncu/f_bicubic -f $B > gpurun_out/f_bicubic_ncu.log 2>&1
echo "rc=$?"
ncu -i /tmp/ncu/f_bicubic.ncu-rep --page raw --csv > gpurun_out/f_bicubic_raw.csv 2>/dev/null
ncu -i /tmp/ncu/f_bicubic.ncu-rep --page source --csv 2>/dev/null | gzip -9 > gpurun_out/f_bicubic_source.csv.gz
ls -la gpurun_out/f_*
