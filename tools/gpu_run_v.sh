#!/bin/bash
# ncu --set full: K1f (float32 bilinear) and K4m akima at its default occupancy
mkdir -p gpurun_out /tmp/ncu
NCU="ncu --set full --clock-control none --import-source on"
B="python bench.py --no-cpu --no-secondary --steps 2 --warmup 1"
for w in bilinear_f32:bilinear2d_tolf_kernel:2e8 akima2d:winmodel_kernel:2e7; do
  name=$(echo $w | cut -d: -f1); kern=$(echo $w | cut -d: -f2); pts=$(echo $w | cut -d: -f3)
  timeout 300 $B --workload $name --points $pts > gpurun_out/v_$name.json 2> gpurun_out/v_$name.err && \
  timeout 600 $NCU -k regex:$kern -s 3 -c 1 -o /tmp/ncu/v_$name -f $B --workload $name --points $pts > gpurun_out/v_${name}_ncu.log 2>&1
  echo "$name rc=$?"
  ncu -i /tmp/ncu/v_$name.ncu-rep --page raw --csv > gpurun_out/v_${name}_raw.csv 2>/dev/null
  python tools/ncu_summary.py gpurun_out/v_${name}_raw.csv $pts > gpurun_out/v_${name}_ncu_summary.txt 2>&1
  tail -1 gpurun_out/v_${name}_ncu_summary.txt
done
python __graft_entry__.py smoke 2>&1 | tail -1
