#!/bin/bash
# round 2, GPU call B: ncu captures (K1t variants, K2/K3 fast kernels), launch list, thread tests
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -s > gpurun_out/b_multi.log 2>&1; echo "multi rc=$?"; tail -4 gpurun_out/b_multi.log
B="python bench.py --no-cpu --no-secondary --steps 2 --warmup 1"
for v in 23 24 44; do
  PYI_TOL_VARIANT=$v timeout 300 $B --points 2e8 > gpurun_out/b_k1_$v.json 2> gpurun_out/b_k1_$v.err && \
  PYI_TOL_VARIANT=$v timeout 600 $NCU -k regex:bilinear2d_tol -s 3 -c 1 -o gpurun_out/b_k1_$v -f $B --points 2e8 > gpurun_out/b_k1_${v}_ncu.log 2>&1
  echo "k1 $v rc=$?"
done
for w in trivariate quadrivariate bilinear_irrlat bicubic; do
  k=bilinear_nd_fast; [ $w = bilinear_irrlat ] && k=bilinear2d_tol; [ $w = bicubic ] && k=bicubicw
  timeout 600 $B --workload $w --points 5e7 > gpurun_out/b_$w.json 2> gpurun_out/b_$w.err && \
  timeout 900 $NCU -k regex:$k -s 3 -c 1 -o gpurun_out/b_$w -f $B --workload $w --points 5e7 > gpurun_out/b_${w}_ncu.log 2>&1
  echo "$w rc=$?"
done
# launch list of the default command (all workloads, reduced point counts)
timeout 900 python bench.py --no-cpu --steps 2 --warmup 1 --points 2e8 > gpurun_out/b_all.json 2> gpurun_out/b_all.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/b_launches.csv \
  python bench.py --no-cpu --steps 2 --warmup 1 --points 2e8 > gpurun_out/b_launches.log 2>&1
echo "launch list rc=$?"
ls -la gpurun_out | head -40
