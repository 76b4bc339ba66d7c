#!/bin/bash
# round 2, GPU call C: full GPU tests on the re-entrant host path, ncu captures reduced to CSV on the box
mkdir -p gpurun_out /tmp/ncu
NCU="ncu --set full --clock-control none --import-source on"
timeout 900 python -m pytest tests -m gpu -x -q -s > gpurun_out/c_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/c_tests.log
B="python bench.py --no-cpu --no-secondary --steps 2 --warmup 1"
extract() {  # $1 = report stem
  ncu -i /tmp/ncu/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null
  ncu -i /tmp/ncu/$1.ncu-rep --page details > gpurun_out/$1_details.txt 2>/dev/null
  ncu -i /tmp/ncu/$1.ncu-rep --page source --csv 2>/dev/null | gzip -9 > gpurun_out/$1_source.csv.gz
}
for v in 23 24 44; do
  PYI_TOL_VARIANT=$v timeout 300 $B --points 2e8 > gpurun_out/c_k1_$v.json 2> gpurun_out/c_k1_$v.err && \
  PYI_TOL_VARIANT=$v timeout 600 $NCU -k regex:bilinear2d_tol -s 3 -c 1 -o /tmp/ncu/c_k1_$v -f $B --points 2e8 > gpurun_out/c_k1_${v}_ncu.log 2>&1
  echo "k1 $v rc=$?"; extract c_k1_$v
done
for w in trivariate quadrivariate bilinear_irrlat bicubic; do
  k=bilinear_nd_fast; [ $w = bilinear_irrlat ] && k=bilinear2d_tol; [ $w = bicubic ] && k=bicubicw
  timeout 600 $B --workload $w --points 5e7 > gpurun_out/c_$w.json 2> gpurun_out/c_$w.err && \
  timeout 900 $NCU -k regex:$k -s 3 -c 1 -o /tmp/ncu/c_$w -f $B --workload $w --points 5e7 > gpurun_out/c_${w}_ncu.log 2>&1
  echo "$w rc=$?"; extract c_$w
done
timeout 900 python bench.py --no-cpu --steps 2 --warmup 1 --points 2e8 > gpurun_out/c_all.json 2> gpurun_out/c_all.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/c_launches.csv \
  python bench.py --no-cpu --steps 2 --warmup 1 --points 2e8 > gpurun_out/c_launches.log 2>&1
echo "launch list rc=$?"
du -sh gpurun_out; ls -la gpurun_out | head -60
