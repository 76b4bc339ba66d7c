#!/bin/bash
# final ncu captures of round 2 (reduced to CSV / text on the box)
mkdir -p gpurun_out /tmp/ncu
NCU="ncu --set full --clock-control none --import-source on"
B="python bench.py --no-cpu --no-secondary --steps 2 --warmup 1"
extract() {
  ncu -i /tmp/ncu/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null
  ncu -i /tmp/ncu/$1.ncu-rep --page source --csv 2>/dev/null | gzip -9 > gpurun_out/$1_source.csv.gz
}
for w in bilinear:2e8:bilinear2d_tol bicubic:5e7:bicubic_sep trivariate_f64:5e7:bilinear_nd_fast idw:5e7:bilinear2d_fast bilinear_irrlat:2e8:bilinear2d_tol; do
  IFS=: read name pts k <<< "$w"
  timeout 600 $B --workload $name --points $pts > gpurun_out/j_$name.json 2> gpurun_out/j_$name.err && \
  timeout 900 $NCU -k regex:$k -s 3 -c 1 -o /tmp/ncu/j_$name -f $B --workload $name --points $pts > gpurun_out/j_${name}_ncu.log 2>&1
  echo "$name rc=$?"; extract j_$name
done
timeout 900 python bench.py --no-cpu --steps 2 --warmup 1 --points 2e8 > gpurun_out/j_all.json 2> gpurun_out/j_all.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/j_launches.csv \
  python bench.py --no-cpu --steps 2 --warmup 1 --points 2e8 > gpurun_out/j_launches.log 2>&1
echo "launch list rc=$?"; du -sh gpurun_out
