#!/bin/bash
# ncu --set full of K4m (akima, 2-D) and of K4s with the c_spline cubics, then the driver's round-end
# commands on one GPU (tools/gpu_run_k.sh)
mkdir -p gpurun_out /tmp/ncu
NCU="ncu --set full --clock-control none --import-source on"
B="python bench.py --no-cpu --no-secondary --steps 2 --warmup 1"
for w in akima2d:winmodel_kernel c_spline2d:bicubic_sep_kernel polynomial2d:winmodel_kernel; do
  name=${w%%:*}; kern=${w##*:}
  timeout 300 $B --workload $name --points 2e7 > gpurun_out/q_$name.json 2> gpurun_out/q_$name.err && \
  timeout 600 $NCU -k regex:$kern -s 3 -c 1 -o /tmp/ncu/q_$name -f $B --workload $name --points 2e7 > gpurun_out/q_${name}_ncu.log 2>&1
  echo "$name rc=$?"
  ncu -i /tmp/ncu/q_$name.ncu-rep --page raw --csv > gpurun_out/q_${name}_raw.csv 2>/dev/null
  python tools/ncu_summary.py gpurun_out/q_${name}_raw.csv 2e7 > gpurun_out/q_${name}_ncu_summary.txt 2>&1
done
bash tools/gpu_run_k.sh 1
