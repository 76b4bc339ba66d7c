#!/bin/bash
# K4m: resident blocks per SM (PYINTERP_B200_M_MINB = 5 | 8 | 10), 2e7 points, 2-D models
mkdir -p gpurun_out
for w in akima2d steffen2d polynomial2d; do
  for b in 5 8 10; do
    PYINTERP_B200_M_MINB=$b timeout 200 python bench.py --workload $w --points 2e7 --steps 5 --warmup 3 --no-cpu --no-secondary \
      > gpurun_out/r_${w}_$b.json 2> gpurun_out/r_${w}_$b.err
    echo "$w minb=$b $(python -c "
import json
d=json.loads(open('gpurun_out/r_${w}_$b.json').read().strip().splitlines()[-1])
print(round(d['value']), 'Mpts/s')" 2>&1 | tail -1)"
  done
done
