#!/bin/bash
# K4m (pyi_winmodels.cuh): the windowed-model tests, then short bench runs of the new workloads,
# each also with the generic kernel (PYINTERP_B200_BICUBIC=exact) for the before / after figure
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_windowed_nd.py tests/test_gpu_fuzz_shapes.py tests/test_gpu_parity.py -x -q -m gpu \
  -k "models or stencil or random_windowed or windowed2d" > gpurun_out/p_tests.log 2>&1
echo "tests rc=$?"; tail -n 4 gpurun_out/p_tests.log
for w in ${WORKLOADS:-akima2d steffen2d c_spline3d polynomial2d}; do
  for mode in tolerance exact; do
    pts=2e7; [ $mode = exact ] && pts=4e6
    PYINTERP_B200_BICUBIC=$mode timeout 300 python bench.py --workload $w --points $pts --steps 5 --warmup 3 --no-cpu \
      > gpurun_out/p_${w}_$mode.json 2> gpurun_out/p_${w}_$mode.err
    echo "$w $mode rc=$? $(python -c "
import json,sys
d=json.loads(open('gpurun_out/p_${w}_$mode.json').read().strip().splitlines()[-1])
print(round(d['value']), 'Mpts/s', d['config']['parity'])" 2>&1 | tail -1)"
  done
done
